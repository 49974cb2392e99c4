// C ABI entry points of libpuschdec_cuda.so (declared in include/pusch_dec_cuda.h). See pdc_runtime_internal.cuh for
// the context layout and the batch planner.
#include "pdc_runtime_internal.cuh"
#include <algorithm>
#include <cstdlib>
#include <thread>

static pdc_status alloc_ticket(pdc_ctx* ctx, ticket_t& t)
{
  const pdc_config& c = ctx->cfg;
  CK(cudaStreamCreateWithFlags(&t.stream, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&t.done, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&t.staged, cudaEventDisableTiming));
  size_t bits_bytes = (size_t)c.max_batch_cbs * (PDC_MAX_K / 8);
  // Transport-block descriptors / results travel in the same copies as the code-block ones: they are laid out right behind
  // the n code-block entries of a submission, so each array has room for both.
  const size_t desc_bytes = (sizeof(pdc_dev_cb) + sizeof(pdc_dev_tb)) * c.max_batch_cbs;
  const size_t res_bytes  = (sizeof(pdc_cb_result) + sizeof(pdc_tb_result)) * c.max_batch_cbs;
  CK(cudaHostAlloc((void**)&t.h_cbs, desc_bytes, cudaHostAllocDefault));
  CK(cudaHostAlloc((void**)&t.h_order, sizeof(uint32_t) * c.max_batch_cbs, cudaHostAllocDefault));
  CK(cudaHostAlloc((void**)&t.h_llrs, c.max_batch_llrs, cudaHostAllocDefault));
  CK(cudaHostAlloc((void**)&t.h_bits, bits_bytes, cudaHostAllocDefault));
  CK(cudaHostAlloc((void**)&t.h_res, res_bytes, cudaHostAllocDefault));
  CK(cudaMalloc((void**)&t.d_cbs, desc_bytes));
  CK(cudaMalloc((void**)&t.d_order, sizeof(uint32_t) * c.max_batch_cbs));
  CK(cudaMalloc((void**)&t.d_llrs, c.max_batch_llrs));
  CK(cudaMalloc((void**)&t.d_bits, bits_bytes));
  CK(cudaMalloc((void**)&t.d_res, res_bytes));
  // Transport-block path: a TB is never larger than its code blocks together, and there are at most as many TBs as CBs.
  CK(cudaHostAlloc((void**)&t.h_tbout, bits_bytes, cudaHostAllocDefault));
  CK(cudaMalloc((void**)&t.d_terms, sizeof(uint32_t) * c.max_batch_cbs));
  CK(cudaMalloc((void**)&t.d_tbout, bits_bytes));
  return PDC_OK;
}

static void free_ticket(ticket_t& t)
{
  if (t.stream) {
    cudaStreamSynchronize(t.stream);
  }
  cudaFreeHost(t.h_cbs), cudaFreeHost(t.h_order), cudaFreeHost(t.h_llrs), cudaFreeHost(t.h_bits);
  cudaFreeHost(t.h_res), cudaFreeHost(t.h_soft);
  cudaFree(t.d_cbs), cudaFree(t.d_order), cudaFree(t.d_llrs), cudaFree(t.d_bits), cudaFree(t.d_res);
  cudaFree(t.d_soft);
  cudaFreeHost(t.h_tbout);
  cudaFree(t.d_terms), cudaFree(t.d_tbout);
  for (cudaEvent_t e : t.tr) {
    if (e) {
      cudaEventDestroy(e);
    }
  }
  if (t.done) {
    cudaEventDestroy(t.done);
  }
  if (t.staged) {
    cudaEventDestroy(t.staged);
  }
  if (t.stream) {
    cudaStreamDestroy(t.stream);
  }
}

// Grows the (rarely used) host-exchanged soft-buffer staging of a ticket.
static pdc_status ensure_soft(pdc_ctx* ctx, ticket_t& t, size_t bytes)
{
  if (bytes <= t.soft_cap) {
    return PDC_OK;
  }
  cudaFreeHost(t.h_soft), cudaFree(t.d_soft);
  t.h_soft = nullptr, t.d_soft = nullptr, t.soft_cap = 0;
  CK(cudaHostAlloc((void**)&t.h_soft, bytes, cudaHostAllocDefault));
  CK(cudaMalloc((void**)&t.d_soft, bytes));
  t.soft_cap = bytes;
  return PDC_OK;
}

extern "C" {

int pdc_abi_version(void)
{
  return PDC_ABI_VERSION;
}

const char* pdc_last_error(const pdc_ctx* ctx)
{
  // A copy owned by the calling thread: other threads may fail (and rewrite the context's string) at any time.
  static thread_local std::string mine;
  if (ctx) {
    pdc_ctx*                    c = const_cast<pdc_ctx*>(ctx);
    std::lock_guard<std::mutex> lock(c->err_mtx);
    mine = c->err;
  } else {
    std::lock_guard<std::mutex> lock(g_static_err_mtx);
    mine = g_static_err;
  }
  return mine.c_str();
}

void* pdc_host_alloc(size_t bytes)
{
  void* p = nullptr;
  if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}

void pdc_host_free(void* p)
{
  if (p) {
    cudaFreeHost(p);
  }
}

pdc_status pdc_set_zero_copy(pdc_ctx* ctx, int enable)
{
  if (!ctx) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  ctx->zero_copy = enable != 0;
  return PDC_OK;
}

pdc_status pdc_create(const pdc_config* cfg, pdc_ctx** out)
{
  if (!cfg || !out || cfg->max_batch_cbs == 0 || cfg->max_batch_llrs == 0 || cfg->nof_tickets == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  *out      = nullptr;
  int ndev  = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
    std::lock_guard<std::mutex> lock(g_static_err_mtx);
    g_static_err = "no CUDA device: this library has no CPU fallback";
    return PDC_ERR_NO_DEVICE;
  }
  pdc_ctx* ctx = new pdc_ctx;
  ctx->cfg     = *cfg;
  ctx->harq_nz.assign(cfg->harq_slots, 0); // the arena is created zeroed
  ctx->harq_released.assign(cfg->harq_slots, 0);
  if (const char* tr = std::getenv("PDC_TRACE")) {
    ctx->trace = tr[0] == '1';
  }
  auto fail    = [&](pdc_status s) {
    {
      std::lock_guard<std::mutex> lock(g_static_err_mtx);
      g_static_err = ctx->err;
    }
    pdc_destroy(ctx);
    return s;
  };
  if (!cuda_ok(ctx, cudaSetDevice(cfg->device), "cudaSetDevice")) {
    return fail(PDC_ERR_CUDA);
  }
  std::vector<uint32_t> graphs, crc, crc_bit;
  pdc_bg_rows           rows;
  build_graph_tables(graphs, rows);
  build_crc_tables(crc);
  build_crc_bit_tables(crc, crc_bit);
  if (!cuda_ok(ctx, pdc_kernels_init_device(&rows), "pdc_kernels_init_device") ||
      !cuda_ok(ctx, pdc_demod_init_device(), "pdc_demod_init_device") ||
      !cuda_ok(ctx, pdc_enc_init_device(), "pdc_enc_init_device") ||
      !cuda_ok(ctx, cudaMalloc((void**)&ctx->d_graphs, graphs.size() * 4), "cudaMalloc graphs") ||
      !cuda_ok(ctx, cudaMemcpy(ctx->d_graphs, graphs.data(), graphs.size() * 4, cudaMemcpyHostToDevice), "copy") ||
      !cuda_ok(ctx, cudaMalloc((void**)&ctx->d_crc, crc.size() * 4), "cudaMalloc crc") ||
      !cuda_ok(ctx, cudaMemcpy(ctx->d_crc, crc.data(), crc.size() * 4, cudaMemcpyHostToDevice), "copy") ||
      !cuda_ok(ctx, cudaMalloc((void**)&ctx->d_crc_bit, crc_bit.size() * 4), "cudaMalloc crc bit table") ||
      !cuda_ok(ctx, cudaMemcpy(ctx->d_crc_bit, crc_bit.data(), crc_bit.size() * 4, cudaMemcpyHostToDevice), "copy")) {
    return fail(PDC_ERR_CUDA);
  }
  if (cfg->harq_slots != 0) {
    size_t bytes = (size_t)cfg->harq_slots * PDC_MAX_N;
    size_t bits  = (size_t)cfg->harq_slots * PDC_SLOT_BITS;
    if (!cuda_ok(ctx, cudaMalloc((void**)&ctx->d_harq, bytes), "cudaMalloc harq arena") ||
        !cuda_ok(ctx, cudaMemset(ctx->d_harq, 0, bytes), "cudaMemset harq arena") ||
        !cuda_ok(ctx, cudaMalloc((void**)&ctx->d_harq_bits, bits), "cudaMalloc harq bits") ||
        !cuda_ok(ctx, cudaMemset(ctx->d_harq_bits, 0, bits), "cudaMemset harq bits")) {
      return fail(PDC_ERR_CUDA);
    }
  }
  if (!cuda_ok(ctx, cudaEventCreateWithFlags(&ctx->side_fork, cudaEventDisableTiming), "cudaEventCreate")) {
    return fail(PDC_ERR_CUDA);
  }
  for (int i = 0; i != pdc_ctx::NOF_SIDE_STREAMS; ++i) {
    if (!cuda_ok(ctx, cudaStreamCreateWithFlags(&ctx->side_stream[i], cudaStreamNonBlocking), "cudaStreamCreate") ||
        !cuda_ok(ctx, cudaEventCreateWithFlags(&ctx->side_join[i], cudaEventDisableTiming), "cudaEventCreate")) {
      return fail(PDC_ERR_CUDA);
    }
  }
  ctx->tickets.resize(cfg->nof_tickets);
  for (ticket_t& t : ctx->tickets) {
    if (alloc_ticket(ctx, t) != PDC_OK) {
      return fail(PDC_ERR_CUDA);
    }
  }
  *out = ctx;
  return PDC_OK;
}

void pdc_destroy(pdc_ctx* ctx)
{
  if (!ctx) {
    return;
  }
  cudaSetDevice(ctx->cfg.device);
  if (ctx->trace && ctx->trace_ms[4] > 0) {
    const double k = 1e3 / ctx->trace_ms[4];
    std::fprintf(stderr, "pdc trace: %.0f submissions; per submission: descriptors H2D %.1f us, kernels %.1f us, D2H %.1f us; "
                         "submit returned -> completion seen %.1f us\n",
                 ctx->trace_ms[4], ctx->trace_ms[0] * k, ctx->trace_ms[1] * k, ctx->trace_ms[2] * k, ctx->trace_ms[3] * k);
  }
  for (ticket_t& t : ctx->tickets) {
    free_ticket(t);
  }
  cudaFree(ctx->d_graphs), cudaFree(ctx->d_crc), cudaFree(ctx->d_crc_bit), cudaFree(ctx->d_harq);
  for (int i = 0; i != pdc_ctx::NOF_SIDE_STREAMS; ++i) {
    if (ctx->side_stream[i]) {
      cudaStreamSynchronize(ctx->side_stream[i]);
      cudaStreamDestroy(ctx->side_stream[i]);
      cudaEventDestroy(ctx->side_join[i]);
    }
  }
  if (ctx->side_fork) {
    cudaEventDestroy(ctx->side_fork);
  }
  cudaFree(ctx->d_harq_bits);
  for (pdc_ctx::demod_slot& d : ctx->demod_ring) {
    cudaFree(d.d_desc);
    cudaFreeHost(d.h_desc);
    if (d.done) {
      cudaEventDestroy(d.done);
    }
  }
  for (pdc_ctx::demod_slot& d : ctx->demux_ring) {
    cudaFree(d.d_desc);
    cudaFreeHost(d.h_desc);
    if (d.done) {
      cudaEventDestroy(d.done);
    }
  }
  cudaFree(ctx->d_demux_in), cudaFree(ctx->d_demux_out);
  for (pdc_ctx::demod_slot& d : ctx->enc_ring) {
    cudaFree(d.d_desc);
    if (d.done) {
      cudaEventDestroy(d.done);
    }
  }
  cudaFree(ctx->d_enc_in), cudaFree(ctx->d_enc_out);
  cudaFree(ctx->d_demod_in), cudaFree(ctx->d_demod_out);
  if (ctx->demod_stream) {
    cudaStreamDestroy(ctx->demod_stream);
  }
  delete ctx;
}

// Takes a free ticket (marks it in use) or returns nullptr. The planning and the CUDA calls of a submission then run
// WITHOUT the context lock: tickets own their stream and staging, so several host threads submit concurrently.
static ticket_t* reserve_ticket(pdc_ctx* ctx)
{
  std::lock_guard<std::mutex> lock(ctx->mtx);
  // Round robin: a ticket that pdc_decode_device borrowed is synchronised with before it is used again, so the least
  // recently used one is the one whose work is most likely over.
  const size_t nt = ctx->tickets.size();
  for (size_t k = 0; k != nt; ++k) {
    ticket_t& c = ctx->tickets[(ctx->next_ticket + k) % nt];
    if (!c.in_use) {
      c.in_use         = true;
      c.id             = ctx->next_id++;
      ctx->next_ticket = (unsigned)((ctx->next_ticket + k + 1) % nt);
      return &c;
    }
  }
  return nullptr;
}

static pdc_status release_ticket(pdc_ctx* ctx, ticket_t* t, pdc_status st)
{
  std::lock_guard<std::mutex> lock(ctx->mtx);
  t->in_use = false;
  return st;
}

// Descriptor staging rings (LLR producers, demultiplexer, PDSCH encoder): when one entry is too small, every entry of the ring
// grows, so a caller that repeats the same submission pays for the allocations (which synchronise the device) once, not once per
// entry. Entries in flight are waited for before their buffers go away.
static pdc_status ring_reserve(pdc_ctx* ctx, pdc_ctx::demod_slot (&ring)[8], size_t bytes, bool pinned_image)
{
  for (pdc_ctx::demod_slot& s : ring) {
    if (s.cap >= bytes && (!pinned_image || s.h_cap >= bytes)) {
      continue;
    }
    if (s.used && s.done) {
      CK(cudaEventSynchronize(s.done));
    }
    if (s.cap < bytes) {
      cudaFree(s.d_desc);
      s.d_desc = nullptr, s.cap = 0;
      CK(cudaMalloc(&s.d_desc, 2 * bytes));
      s.cap = 2 * bytes;
    }
    if (pinned_image && s.h_cap < bytes) {
      cudaFreeHost(s.h_desc);
      s.h_desc = nullptr, s.h_cap = 0;
      CK(cudaHostAlloc(&s.h_desc, 2 * bytes, cudaHostAllocDefault));
      s.h_cap = 2 * bytes;
    }
  }
  return PDC_OK;
}

#define CKT(call)                                                                                                      \
  do {                                                                                                                 \
    if (!cuda_ok(ctx, (call), #call)) {                                                                                \
      return release_ticket(ctx, t, PDC_ERR_CUDA);                                                                     \
    }                                                                                                                  \
  } while (0)

// Arena bookkeeping of a planned batch (batch_buffers): where plan_batch keeps its scratch, and the clearing of released slots
// that the batch touches without rewriting them (stream-ordered, before the kernels).
static void arena_tracking(pdc_ctx* ctx, batch_buffers& buf, int32_t* nz_scratch, int32_t* zero_slots, uint32_t* nof_zero_slots)
{
  *nof_zero_slots    = 0;
  buf.harq_nz        = ctx->harq_nz.data();
  buf.nz_scratch     = nz_scratch;
  buf.harq_released  = ctx->harq_released.data();
  buf.zero_slots     = zero_slots;
  buf.nof_zero_slots = nof_zero_slots;
}
static cudaError_t clear_released_slots(pdc_ctx* ctx, const int32_t* zero_slots, uint32_t nof_zero_slots, cudaStream_t stream)
{
  for (uint32_t k = 0; k != nof_zero_slots; ++k) {
    cudaError_t e = cudaMemsetAsync(ctx->d_harq + (size_t)zero_slots[k] * PDC_MAX_N, 0, PDC_MAX_N, stream);
    if (e != cudaSuccess) {
      return e;
    }
  }
  return cudaSuccess;
}

// pdc_submit (tbs == nullptr) and pdc_submit_tbs.
// stage != nullptr (pdc_submit_staged): the LLRs already lie in the stage's device arena (llr_offset relative to its base).
static pdc_status submit_common(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs, uint32_t ntb,
                                const int8_t* llrs, size_t llr_bytes, const int8_t* soft_in, size_t soft_bytes,
                                pdc_ticket* ticket, pdc_stage* stage = nullptr)
{
  if (!ctx || !descs || !ticket || n == 0 || (!llrs && llr_bytes != 0 && !stage) || (tbs && ntb == 0)) {
    return PDC_ERR_INVALID_ARG;
  }
  if (n > ctx->cfg.max_batch_cbs || (!stage && llr_bytes > ctx->cfg.max_batch_llrs) || ntb > ctx->cfg.max_batch_cbs) {
    return PDC_ERR_CAPACITY;
  }
  nvtx_scope range(tbs ? "pdc_submit_tbs" : "pdc_submit");
  ticket_t*  t = reserve_ticket(ctx);
  if (!t) {
    return PDC_ERR_CAPACITY;
  }
  CKT(cudaSetDevice(ctx->cfg.device));
  if (t->borrowed) {
    // pdc_decode_device's descriptor copy out of this pinned staging has to be over; its kernels, which read the device
    // staging, are ordered before this ticket's stream through t->done (cudaStreamWaitEvent in decode_device_common).
    CKT(cudaEventSynchronize(t->staged));
    t->borrowed = false;
  }
  batch_buffers buf;
  // Staged code-block submissions (the accelerator adapters: small batches, latency matters): the kernels store the hard bits and
  // the results straight into the page-locked host buffers (device-addressable), so no device-to-host copy follows them.
  const bool host_out = stage != nullptr && !tbs;
  buf.d_llrs = stage ? stage->d_base : t->d_llrs, buf.d_soft = t->d_soft;
  buf.d_bits = host_out ? t->h_bits : t->d_bits, buf.d_res = host_out ? t->h_res : t->d_res;
  buf.d_harq = ctx->d_harq, buf.harq_slots = ctx->cfg.harq_slots, buf.llr_bytes = stage ? stage->bytes : llr_bytes;
  buf.d_slot_bits = tbs ? ctx->d_harq_bits : nullptr;
  batch_plan bp;
  t->lst.resize(n);
  t->nz.resize(n);
  t->zero_slots.resize(n);
  arena_tracking(ctx, buf, t->nz.data(), t->zero_slots.data(), &t->nof_zero_slots);
  pdc_status st = plan_batch(descs, n, buf, t->h_cbs, t->h_order, t->lst.data(), bp);
  if (st != PDC_OK) {
    return release_ticket(ctx, t, st);
  }
  if (bp.bits_end > (size_t)ctx->cfg.max_batch_cbs * (PDC_MAX_K / 8)) {
    return release_ticket(ctx, t, PDC_ERR_CAPACITY);
  }
  tb_plan tp;
  if (tbs) {
    t->h_tbs   = reinterpret_cast<pdc_dev_tb*>(t->h_cbs + n), t->d_tbs = reinterpret_cast<pdc_dev_tb*>(t->d_cbs + n);
    t->h_tbres = reinterpret_cast<pdc_tb_result*>(t->h_res + n), t->d_tbres = reinterpret_cast<pdc_tb_result*>(t->d_res + n);
    st = plan_tbs(tbs, ntb, descs, n, t->d_tbout, t->d_tbres, t->d_terms, t->h_tbs, tp);
    if (st == PDC_OK && tp.bytes_end > (size_t)ctx->cfg.max_batch_cbs * (PDC_MAX_K / 8)) {
      st = PDC_ERR_CAPACITY;
    }
    if (st != PDC_OK) {
      return release_ticket(ctx, t, st);
    }
  }
  if (bp.any_host_soft) {
    if (soft_in && soft_bytes < bp.soft_end) {
      return release_ticket(ctx, t, PDC_ERR_INVALID_ARG);
    }
    if (bp.soft_end > t->soft_cap) {
      // The staging grows (rare): the device pointers planned above are stale, plan again on the new buffers.
      st = ensure_soft(ctx, *t, bp.soft_end);
      if (st != PDC_OK) {
        return release_ticket(ctx, t, st);
      }
      buf.d_soft = t->d_soft;
      // The arena bookkeeping of the first pass stands (planning again from the state it left gives the same decisions);
      // only the list of released slots to clear must not be started over.
      buf.zero_slots = nullptr, buf.nof_zero_slots = nullptr;
      batch_plan again;
      st = plan_batch(descs, n, buf, t->h_cbs, t->h_order, t->lst.data(), again);
      if (st != PDC_OK) {
        return release_ticket(ctx, t, st);
      }
      bp = again;
    }
  }
  CKT(clear_released_slots(ctx, t->zero_slots.data(), t->nof_zero_slots, t->stream));
  // Stage and go. With pdc_set_zero_copy, LLRs that already lie in pinned (or registered) memory are sent from where they
  // lie: the caller then keeps them untouched until the ticket has completed.
  const int8_t* src    = llrs;
  bool          pinned = false;
  if (stage) {
    // Everything pushed so far has to be on the device before the kernels start: one event behind the copy stream's tail.
    std::lock_guard<std::mutex> slock(stage->mtx);
    CKT(cudaEventRecord(stage->pushed, stage->copy));
    CKT(cudaStreamWaitEvent(t->stream, stage->pushed, 0));
    llr_bytes = 0;
  }
  if (ctx->trace) {
    for (cudaEvent_t& e : t->tr) {
      if (!e) {
        CKT(cudaEventCreate(&e));
      }
    }
    CKT(cudaEventRecord(t->tr[0], t->stream));
  }
  if (ctx->zero_copy && llr_bytes != 0) {
    cudaPointerAttributes attr;
    pinned = (cudaPointerGetAttributes(&attr, llrs) == cudaSuccess) && (attr.type == cudaMemoryTypeHost);
    cudaGetLastError(); // pageable pointers make the query set an error on some runtimes
  }
  if (!pinned && llr_bytes != 0) {
    memcpy(t->h_llrs, llrs, llr_bytes);
    src = t->h_llrs;
  }
  if (llr_bytes != 0) {
    CKT(cudaMemcpyAsync(t->d_llrs, src, llr_bytes, cudaMemcpyHostToDevice, t->stream));
  }
  if (host_out) {
    // Nothing to upload: each thread block reads its own descriptor from the page-locked image (one 56-byte read over the
    // link per code block, in parallel), see launch_batch below.
  } else if (stage) {
    // Staged submissions run while other regions of the arena are travelling: keep the small copies off the copy engine.
    CKT(pdc_launch_upload(t->d_cbs, t->h_cbs, sizeof(pdc_dev_cb) * n + sizeof(pdc_dev_tb) * ntb, t->d_order, t->h_order,
                          bp.identity ? 0 : sizeof(uint32_t) * n, t->stream));
    ctx->launches += 1;
  } else {
    CKT(cudaMemcpyAsync(t->d_cbs, t->h_cbs, sizeof(pdc_dev_cb) * n + sizeof(pdc_dev_tb) * ntb, cudaMemcpyHostToDevice,
                        t->stream));
    if (!bp.identity) {
      CKT(cudaMemcpyAsync(t->d_order, t->h_order, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, t->stream));
    }
  }
  if (bp.any_host_soft) {
    if (soft_in) {
      memcpy(t->h_soft, soft_in, bp.soft_end);
      CKT(cudaMemcpyAsync(t->d_soft, t->h_soft, bp.soft_end, cudaMemcpyHostToDevice, t->stream));
    } else {
      CKT(cudaMemsetAsync(t->d_soft, 0, bp.soft_end, t->stream));
    }
  }
  if (ctx->trace) {
    CKT(cudaEventRecord(t->tr[1], t->stream));
  }
  st = host_out ? launch_batch(ctx, bp, t->h_cbs, t->h_order, t->stream) : launch_batch(ctx, bp, t->d_cbs, t->d_order, t->stream);
  if (st != PDC_OK) {
    return release_ticket(ctx, t, st);
  }
  if (ctx->trace) {
    CKT(cudaEventRecord(t->tr[2], t->stream));
  }
  if (tbs) {
    // Assembly + TB CRC on the device; the payloads and the two result arrays (one copy) come back, no per-CB bits.
    CKT(pdc_launch_tb(t->d_tbs, t->d_cbs, ctx->d_crc, (int)ntb, (int)tp.max_cbs, t->stream));
    ctx->launches += 2;
    CKT(cudaMemcpyAsync(t->h_tbout, t->d_tbout, tp.bytes_end, cudaMemcpyDeviceToHost, t->stream));
  } else if (!host_out) {
    CKT(cudaMemcpyAsync(t->h_bits, t->d_bits, bp.bits_end, cudaMemcpyDeviceToHost, t->stream));
  }
  if (!host_out) {
    CKT(cudaMemcpyAsync(t->h_res, t->d_res, sizeof(pdc_cb_result) * n + sizeof(pdc_tb_result) * ntb, cudaMemcpyDeviceToHost,
                        t->stream));
  }
  if (bp.any_host_soft) {
    CKT(cudaMemcpyAsync(t->h_soft, t->d_soft, bp.soft_end, cudaMemcpyDeviceToHost, t->stream));
  }
  if (ctx->trace) {
    CKT(cudaEventRecord(t->tr[3], t->stream));
    t->submitted_at = std::chrono::steady_clock::now();
  }
  CKT(cudaEventRecord(t->done, t->stream));
  t->n          = n;
  t->bits_bytes = bp.bits_end;
  t->bits_dense = bp.bits_dense;
  t->soft_bytes = bp.soft_end;
  t->soft_back  = bp.any_host_soft;
  t->descs.assign(descs, descs + n);
  t->ntb = ntb;
  t->tbs.assign(tbs, tbs + (tbs ? ntb : 0));
  *ticket = t->id;
  return PDC_OK;
}

pdc_status pdc_submit(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const int8_t* llrs, size_t llr_bytes,
                      const int8_t* soft_in, size_t soft_bytes, pdc_ticket* ticket)
{
  if (!llrs) {
    return PDC_ERR_INVALID_ARG;
  }
  return submit_common(ctx, descs, n, nullptr, 0, llrs, llr_bytes, soft_in, soft_bytes, ticket);
}

pdc_status pdc_stage_open(pdc_ctx* ctx, size_t bytes, pdc_stage** out)
{
  if (!ctx || !out || bytes == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  *out = nullptr;
  CK(cudaSetDevice(ctx->cfg.device));
  pdc_stage* s = new pdc_stage;
  s->ctx       = ctx;
  s->bytes     = bytes;
  if (!cuda_ok(ctx, cudaHostAlloc((void**)&s->h_base, bytes, cudaHostAllocDefault), "cudaHostAlloc stage") ||
      !cuda_ok(ctx, cudaMalloc((void**)&s->d_base, bytes), "cudaMalloc stage") ||
      !cuda_ok(ctx, cudaStreamCreateWithFlags(&s->copy, cudaStreamNonBlocking), "cudaStreamCreate stage") ||
      !cuda_ok(ctx, cudaEventCreateWithFlags(&s->pushed, cudaEventDisableTiming), "cudaEventCreate stage")) {
    pdc_stage_close(s);
    return PDC_ERR_CUDA;
  }
  *out = s;
  return PDC_OK;
}

void pdc_stage_close(pdc_stage* s)
{
  if (!s) {
    return;
  }
  cudaSetDevice(s->ctx->cfg.device);
  if (s->copy) {
    cudaStreamSynchronize(s->copy);
    cudaStreamDestroy(s->copy);
  }
  if (s->pushed) {
    cudaEventDestroy(s->pushed);
  }
  cudaFreeHost(s->h_base);
  cudaFree(s->d_base);
  delete s;
}

int8_t* pdc_stage_host(pdc_stage* s)
{
  return s ? s->h_base : nullptr;
}

pdc_status pdc_stage_push(pdc_stage* s, size_t offset, size_t bytes)
{
  if (!s || offset > s->bytes || bytes > s->bytes - offset) {
    return PDC_ERR_INVALID_ARG;
  }
  if (bytes == 0) {
    return PDC_OK;
  }
  pdc_ctx* ctx = s->ctx;
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> lock(s->mtx);
  CK(cudaMemcpyAsync(s->d_base + offset, s->h_base + offset, bytes, cudaMemcpyHostToDevice, s->copy));
  return PDC_OK;
}

pdc_status pdc_submit_staged(pdc_ctx* ctx, pdc_stage* stage, const pdc_cb_desc* descs, uint32_t n, const int8_t* soft_in,
                             size_t soft_bytes, pdc_ticket* ticket)
{
  if (!stage || stage->ctx != ctx) {
    return PDC_ERR_INVALID_ARG;
  }
  return submit_common(ctx, descs, n, nullptr, 0, nullptr, 0, soft_in, soft_bytes, ticket, stage);
}

pdc_status pdc_submit_tbs(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs, uint32_t ntb,
                          const int8_t* llrs, size_t llr_bytes, const int8_t* soft_in, size_t soft_bytes,
                          pdc_ticket* ticket)
{
  if (!tbs || ntb == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  return submit_common(ctx, descs, n, tbs, ntb, llrs, llr_bytes, soft_in, soft_bytes, ticket);
}

static ticket_t* find_ticket(pdc_ctx* ctx, pdc_ticket id)
{
  for (ticket_t& c : ctx->tickets) {
    if (c.in_use && c.id == id) {
      return &c;
    }
  }
  return nullptr;
}

pdc_status pdc_poll(pdc_ctx* ctx, pdc_ticket ticket)
{
  if (!ctx) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  ticket_t*                   t = find_ticket(ctx, ticket);
  if (!t) {
    return PDC_ERR_BAD_TICKET;
  }
  cudaError_t e = cudaEventQuery(t->done);
  if (e == cudaErrorNotReady) {
    return PDC_ERR_NOT_READY;
  }
  CK(e);
  return PDC_OK;
}

// Host-exchanged soft buffers of a finished ticket back to the caller.
static pdc_status copy_soft_back(ticket_t* t, int8_t* soft_out, size_t soft_bytes)
{
  pdc_status st = PDC_OK;
  if (soft_out && t->soft_back) {
    for (uint32_t i = 0; i != t->n; ++i) {
      const pdc_cb_desc& d = t->descs[i];
      if (d.harq_slot >= 0 || (d.flags & (PDC_F_DECODE_ONLY | PDC_F_DECODED))) {
        continue;
      }
      size_t N = pdc_cb_buffer_len(d.bg, d.Z);
      if (d.soft_offset + N > soft_bytes) {
        st = PDC_ERR_INVALID_ARG;
      } else {
        memcpy(soft_out + d.soft_offset, t->h_soft + d.soft_offset, N);
      }
    }
  }
  return st;
}

// Finds the ticket and waits for it (outside the lock: other threads keep submitting).
static pdc_status wait_ticket(pdc_ctx* ctx, pdc_ticket ticket, bool want_tbs, ticket_t*& t)
{
  {
    std::lock_guard<std::mutex> lock(ctx->mtx);
    t = find_ticket(ctx, ticket);
    if (!t || (t->ntb != 0) != want_tbs) {
      return PDC_ERR_BAD_TICKET;
    }
  }
  // Query + yield instead of cudaEventSynchronize: the runtime's wait spins on a core, and the callers of this library are
  // pools of host threads with work of their own (the reference's decoder threads); an idle machine loses nothing.
  cudaError_t e;
  while ((e = cudaEventQuery(t->done)) == cudaErrorNotReady) {
    std::this_thread::yield();
  }
  if (!cuda_ok(ctx, e, "cudaEventQuery")) {
    return release_ticket(ctx, t, PDC_ERR_CUDA);
  }
  if (ctx->trace && t->tr[3]) {
    const double host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t->submitted_at).count();
    float        a = 0, b = 0, c = 0;
    cudaEventElapsedTime(&a, t->tr[0], t->tr[1]), cudaEventElapsedTime(&b, t->tr[1], t->tr[2]);
    cudaEventElapsedTime(&c, t->tr[2], t->tr[3]);
    std::lock_guard<std::mutex> lock(ctx->mtx);
    ctx->trace_ms[0] += a, ctx->trace_ms[1] += b, ctx->trace_ms[2] += c, ctx->trace_ms[3] += host_ms, ctx->trace_ms[4] += 1;
  }
  return PDC_OK;
}

pdc_status pdc_fetch_tbs(pdc_ctx* ctx, pdc_ticket ticket, uint8_t* tb_bytes, size_t tb_bytes_len, pdc_tb_result* tb_res,
                         pdc_cb_result* cb_res, int8_t* soft_out, size_t soft_bytes)
{
  if (!ctx || !tb_res) {
    return PDC_ERR_INVALID_ARG;
  }
  nvtx_scope range("pdc_fetch_tbs");
  ticket_t*  t;
  pdc_status st = wait_ticket(ctx, ticket, true, t);
  if (st != PDC_OK) {
    return st;
  }
  memcpy(tb_res, t->h_tbres, sizeof(pdc_tb_result) * t->ntb);
  if (cb_res) {
    memcpy(cb_res, t->h_res, sizeof(pdc_cb_result) * t->n);
    for (uint32_t i = 0; i != t->n; ++i) {
      if (t->descs[i].flags & PDC_F_DECODED) { // no kernel wrote this entry
        cb_res[i].iterations = -1, cb_res[i].crc_pass = 1, cb_res[i].status = 0, cb_res[i].nof_layers = 0;
      }
    }
  }
  if (tb_bytes) {
    for (uint32_t k = 0; k != t->ntb; ++k) {
      const pdc_tb_desc& tb = t->tbs[k];
      if (!tb_res[k].written) {
        continue; // the caller's transport block buffer stays untouched (pusch_decoder_impl.cpp:425)
      }
      if (tb.tb_offset + tb.tbs_bits / 8 > tb_bytes_len) {
        st = PDC_ERR_INVALID_ARG;
      } else {
        memcpy(tb_bytes + tb.tb_offset, t->h_tbout + tb.tb_offset, tb.tbs_bits / 8);
      }
    }
  }
  pdc_status s2 = copy_soft_back(t, soft_out, soft_bytes);
  return release_ticket(ctx, t, st != PDC_OK ? st : s2);
}

pdc_status pdc_fetch(pdc_ctx* ctx, pdc_ticket ticket, uint8_t* bits, size_t bits_bytes, pdc_cb_result* res,
                     int8_t* soft_out, size_t soft_bytes)
{
  if (!ctx || !res) {
    return PDC_ERR_INVALID_ARG;
  }
  nvtx_scope range("pdc_fetch");
  ticket_t*  t;
  pdc_status st = wait_ticket(ctx, ticket, false, t);
  if (st != PDC_OK) {
    return st;
  }
  memcpy(res, t->h_res, sizeof(pdc_cb_result) * t->n);
  bool all_written = true; // status 2: "not enough soft bits" with a CRC object — the reference leaves the output untouched
  for (uint32_t i = 0; i != t->n; ++i) {
    all_written &= (t->h_res[i].status == 0);
  }
  if (bits && all_written && t->bits_dense && t->bits_bytes <= bits_bytes) {
    memcpy(bits, t->h_bits, t->bits_bytes); // the outputs tile the range: one copy (words are whole: K is a multiple of 8
                                            // only up to the word padding each code block owns anyway)
  } else if (bits) {
    for (uint32_t i = 0; i != t->n; ++i) {
      const pdc_cb_desc& d = t->descs[i];
      if (res[i].status != 0 || (d.flags & PDC_F_DEMATCH_ONLY)) {
        continue;
      }
      size_t nb = ((size_t)pdc_cb_message_bits(d.bg, d.Z) + 7) / 8;
      if (d.out_offset + nb > bits_bytes) {
        st = PDC_ERR_INVALID_ARG;
      } else {
        memcpy(bits + d.out_offset, t->h_bits + d.out_offset, nb);
      }
    }
  }
  pdc_status s2 = copy_soft_back(t, soft_out, soft_bytes);
  return release_ticket(ctx, t, st != PDC_OK ? st : s2);
}

static pdc_status decode_device_common(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs,
                                       uint32_t ntb, const int8_t* d_llrs, int8_t* d_soft, uint8_t* d_bits,
                                       pdc_cb_result* d_res, uint8_t* d_tb_bytes, pdc_tb_result* d_tb_res, void* stream_v)
{
  if (!ctx || !descs || !d_res || n == 0 || (tbs ? (ntb == 0 || !d_tb_bytes || !d_tb_res) : (!d_llrs || !d_bits))) {
    return PDC_ERR_INVALID_ARG;
  }
  if (n > ctx->cfg.max_batch_cbs || ntb > ctx->cfg.max_batch_cbs) {
    return PDC_ERR_CAPACITY;
  }
  nvtx_scope   range(tbs ? "pdc_decode_device_tbs" : "pdc_decode_device");
  cudaStream_t stream = (cudaStream_t)stream_v;
  // The descriptor staging of a free ticket is borrowed for the duration of the launch (stream-ordered).
  ticket_t* t = reserve_ticket(ctx);
  if (!t) {
    return PDC_ERR_CAPACITY;
  }
  CKT(cudaSetDevice(ctx->cfg.device));
  // A previous borrower's asynchronous descriptor copy may still be reading the pinned staging (host side: wait for that
  // copy only), and its kernels may still be reading the device staging from ANOTHER stream (device side: this stream waits
  // for them; a no-op when it is the same stream).
  if (t->borrowed) {
    CKT(cudaEventSynchronize(t->staged));
  }
  CKT(cudaStreamWaitEvent(stream, t->done, 0));
  t->borrowed = true;
  batch_buffers buf;
  buf.d_llrs = d_llrs, buf.d_soft = d_soft, buf.d_bits = d_bits, buf.d_res = d_res;
  buf.d_harq = ctx->d_harq, buf.harq_slots = ctx->cfg.harq_slots;
  buf.d_slot_bits = tbs ? ctx->d_harq_bits : nullptr;
  batch_plan bp;
  t->lst.resize(n);
  t->nz.resize(n);
  t->zero_slots.resize(n);
  arena_tracking(ctx, buf, t->nz.data(), t->zero_slots.data(), &t->nof_zero_slots);
  pdc_status st = plan_batch(descs, n, buf, t->h_cbs, t->h_order, t->lst.data(), bp);
  if (st == PDC_OK && bp.any_host_soft && !d_soft) {
    st = PDC_ERR_INVALID_ARG;
  }
  tb_plan tp;
  if (st == PDC_OK && tbs) {
    t->h_tbs = reinterpret_cast<pdc_dev_tb*>(t->h_cbs + n), t->d_tbs = reinterpret_cast<pdc_dev_tb*>(t->d_cbs + n);
    st       = plan_tbs(tbs, ntb, descs, n, d_tb_bytes, d_tb_res, t->d_terms, t->h_tbs, tp);
  }
  if (st != PDC_OK) {
    return release_ticket(ctx, t, st);
  }
  CKT(clear_released_slots(ctx, t->zero_slots.data(), t->nof_zero_slots, stream));
  CKT(cudaMemcpyAsync(t->d_cbs, t->h_cbs, sizeof(pdc_dev_cb) * n + sizeof(pdc_dev_tb) * ntb, cudaMemcpyHostToDevice, stream));
  if (!bp.identity) {
    CKT(cudaMemcpyAsync(t->d_order, t->h_order, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, stream));
  }
  // The pinned staging is free again once the descriptor copies have been consumed.
  CKT(cudaEventRecord(t->staged, stream));
  st = launch_batch(ctx, bp, t->d_cbs, t->d_order, stream);
  if (st == PDC_OK && tbs) {
    CKT(pdc_launch_tb(t->d_tbs, t->d_cbs, ctx->d_crc, (int)ntb, (int)tp.max_cbs, stream));
    ctx->launches += 2;
  }
  // d_cbs / d_order (and d_tbs / d_terms) of this ticket are read by the launches above: the next user of the ticket, on
  // whatever stream, orders itself after them through the same event, recorded again behind the kernels.
  CKT(cudaEventRecord(t->done, stream));
  CKT(cudaStreamWaitEvent(t->stream, t->done, 0));
  return release_ticket(ctx, t, st);
}

pdc_status pdc_decode_device(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const int8_t* d_llrs, int8_t* d_soft,
                             uint8_t* d_bits, pdc_cb_result* d_res, void* stream)
{
  return decode_device_common(ctx, descs, n, nullptr, 0, d_llrs, d_soft, d_bits, d_res, nullptr, nullptr, stream);
}

pdc_status pdc_decode_device_tbs(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs, uint32_t ntb,
                                 const int8_t* d_llrs, int8_t* d_soft, uint8_t* d_bits, pdc_cb_result* d_res,
                                 uint8_t* d_tb_bytes, pdc_tb_result* d_tb_res, void* stream)
{
  if (!tbs || ntb == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  return decode_device_common(ctx, descs, n, tbs, ntb, d_llrs, d_soft, d_bits, d_res, d_tb_bytes, d_tb_res, stream);
}

// ---- allocation-free variant of pdc_decode_device_tbs ------------------------------------------------------------------------
// The device scratch the transport-block path needs between its kernels — device descriptors of the code blocks and transport
// blocks, the launch order and every code block's share of the TB CRC (pdc_dev_tb.terms, written by pdc_tb_assemble_kernel,
// read by pdc_tb_finish_kernel) — lives in ONE file-scope device array instead of a ticket's cudaMalloc'ed staging. Every GPU
// that loads the library holds these PDC_STATIC_SCRATCH_BYTES until exit. All calls share the array.
#define PDC_STATIC_CBS_OFF 0
#define PDC_STATIC_TBS_OFF (PDC_STATIC_CBS_OFF + sizeof(pdc_dev_cb) * PDC_STATIC_MAX_CBS)
#define PDC_STATIC_ORDER_OFF (PDC_STATIC_TBS_OFF + sizeof(pdc_dev_tb) * PDC_STATIC_MAX_CBS)
#define PDC_STATIC_TERMS_OFF (PDC_STATIC_ORDER_OFF + sizeof(uint32_t) * PDC_STATIC_MAX_CBS)
#define PDC_STATIC_SCRATCH_BYTES (PDC_STATIC_TERMS_OFF + sizeof(uint32_t) * PDC_STATIC_MAX_CBS)
static_assert(PDC_STATIC_TBS_OFF % 16 == 0 && PDC_STATIC_ORDER_OFF % 16 == 0 && PDC_STATIC_TERMS_OFF % 16 == 0, "alignment");

static __device__ __align__(256) unsigned char g_pdc_static_scratch[PDC_STATIC_SCRATCH_BYTES];

// In host code `g_pdc_static_scratch` is only a handle of the symbol, not a device address: the address is asked for once per
// device, with that device current, and only that pointer is ever handed to a kernel or a copy.
#define PDC_STATIC_MAX_DEVICES 64
static unsigned char* g_static_scratch_ptr[PDC_STATIC_MAX_DEVICES];
static std::mutex     g_static_mtx; // serialises the calls on the host: one host image of the descriptors, one device array
static struct {
  pdc_dev_cb cbs[PDC_STATIC_MAX_CBS];
  pdc_dev_tb tbs[PDC_STATIC_MAX_CBS];
  uint32_t   order[PDC_STATIC_MAX_CBS];
  uint8_t    lst[PDC_STATIC_MAX_CBS];
  int32_t    nz[PDC_STATIC_MAX_CBS];
  int32_t    zero_slots[PDC_STATIC_MAX_CBS];
  uint32_t   nof_zero_slots;
} g_static_host; // .bss, nothing is allocated per call

static pdc_status static_scratch_address(pdc_ctx* ctx, unsigned char** out)
{
  const int dev = ctx->cfg.device;
  if (dev < 0 || dev >= PDC_STATIC_MAX_DEVICES) {
    return PDC_ERR_INVALID_ARG;
  }
  CK(cudaSetDevice(dev));
  if (!g_static_scratch_ptr[dev]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_pdc_static_scratch));
    if (!p) {
      std::lock_guard<std::mutex> elock(ctx->err_mtx);
      ctx->err = "cudaGetSymbolAddress returned a null address for the static scratch";
      return PDC_ERR_CUDA;
    }
    g_static_scratch_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_static_scratch_ptr[dev];
  return PDC_OK;
}

pdc_status pdc_decode_device_tbs_static(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs,
                                        uint32_t ntb, const int8_t* d_llrs, int8_t* d_soft, uint8_t* d_bits,
                                        pdc_cb_result* d_res, uint8_t* d_tb_bytes, pdc_tb_result* d_tb_res, void* stream_v)
{
  if (!ctx || !descs || !d_res || n == 0 || !tbs || ntb == 0 || !d_tb_bytes || !d_tb_res) {
    return PDC_ERR_INVALID_ARG;
  }
  if (n > PDC_STATIC_MAX_CBS || ntb > PDC_STATIC_MAX_CBS) {
    std::lock_guard<std::mutex> lock(ctx->err_mtx);
    ctx->err = "pdc_decode_device_tbs_static: more than PDC_STATIC_MAX_CBS (1024) code blocks or transport blocks in one "
               "call; split the batch or use pdc_decode_device_tbs";
    return PDC_ERR_CAPACITY;
  }
  cudaStream_t                stream = (cudaStream_t)stream_v;
  std::lock_guard<std::mutex> lock(g_static_mtx);
  unsigned char*              base = nullptr;
  pdc_status                  st   = static_scratch_address(ctx, &base);
  if (st != PDC_OK) {
    return st;
  }
  pdc_dev_cb* d_cbs   = reinterpret_cast<pdc_dev_cb*>(base + PDC_STATIC_CBS_OFF);
  pdc_dev_tb* d_tbs   = reinterpret_cast<pdc_dev_tb*>(base + PDC_STATIC_TBS_OFF);
  uint32_t*   d_order = reinterpret_cast<uint32_t*>(base + PDC_STATIC_ORDER_OFF);
  uint32_t*   d_terms = reinterpret_cast<uint32_t*>(base + PDC_STATIC_TERMS_OFF);
  batch_buffers buf;
  buf.d_llrs = d_llrs, buf.d_soft = d_soft, buf.d_bits = d_bits, buf.d_res = d_res;
  buf.d_harq = ctx->d_harq, buf.harq_slots = ctx->cfg.harq_slots;
  buf.d_slot_bits = ctx->d_harq_bits;
  batch_plan bp;
  arena_tracking(ctx, buf, g_static_host.nz, g_static_host.zero_slots, &g_static_host.nof_zero_slots);
  st = plan_batch(descs, n, buf, g_static_host.cbs, g_static_host.order, g_static_host.lst, bp);
  if (st == PDC_OK) {
    uint32_t processed = bp.n_lane; // code blocks a kernel reads LLRs for (all but the PDC_F_DECODED ones)
    for (int k = 0; k != PDC_NOF_PACKED_LISTS; ++k) {
      processed += bp.n_h2[k];
    }
    if ((bp.any_host_soft && !d_soft) || (processed != 0 && !d_llrs)) {
      st = PDC_ERR_INVALID_ARG;
    }
  }
  tb_plan tp;
  if (st == PDC_OK) {
    st = plan_tbs(tbs, ntb, descs, n, d_tb_bytes, d_tb_res, d_terms, g_static_host.tbs, tp);
  }
  if (st != PDC_OK) {
    return st;
  }
  // Pageable sources: cudaMemcpyAsync returns once they have been staged, so the host image is free for the next call when
  // the lock is released. The array keeps the previous call's contents: the CRC shares start from zero again.
  CK(clear_released_slots(ctx, g_static_host.zero_slots, g_static_host.nof_zero_slots, stream));
  CK(cudaMemcpyAsync(d_cbs, g_static_host.cbs, sizeof(pdc_dev_cb) * n, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(d_tbs, g_static_host.tbs, sizeof(pdc_dev_tb) * ntb, cudaMemcpyHostToDevice, stream));
  if (!bp.identity) {
    CK(cudaMemcpyAsync(d_order, g_static_host.order, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, stream));
  }
  CK(cudaMemsetAsync(d_terms, 0, sizeof(uint32_t) * n, stream));
  st = launch_batch(ctx, bp, d_cbs, d_order, stream);
  if (st != PDC_OK) {
    return st;
  }
  CK(pdc_launch_tb(d_tbs, d_cbs, ctx->d_crc, (int)ntb, (int)tp.max_cbs, stream));
  ctx->launches += 2;
  return PDC_OK;
}

// ---- LLR producers ---------------------------------------------------------------------------------------------------------
static int demod_bits_per_symbol(unsigned mod)
{
  switch (mod) {
    case PDC_MOD_PI2_BPSK:
    case PDC_MOD_BPSK:
      return 1;
    case PDC_MOD_QPSK:
    case PDC_MOD_QAM16:
    case PDC_MOD_QAM64:
    case PDC_MOD_QAM256:
      return (int)mod;
    default:
      return 0;
  }
}

// Validates the segments against the array sizes (0 = unknown, device entry) and lists the chunks of 1024 symbols.
static pdc_status demod_plan(const pdc_demod_seg* segs, uint32_t n, size_t nof_symbols, size_t llr_bytes,
                             std::vector<uint32_t>& chunks, size_t& max_sym, size_t& max_llr)
{
  chunks.clear();
  max_sym = max_llr = 0;
  for (uint32_t i = 0; i != n; ++i) {
    const pdc_demod_seg& s  = segs[i];
    const int            qm = demod_bits_per_symbol(s.mod);
    if (qm == 0 || (s.flags & ~PDC_DEMOD_F_DESCRAMBLE) != 0 || s.reserved != 0 ||
        ((uint64_t)s.scramble_offset + (uint64_t)s.nof_symbols * qm + 1600u) >= (1ull << 28)) {
      return PDC_ERR_INVALID_ARG;
    }
    const size_t se = (size_t)s.sym_offset + s.nof_symbols, le = (size_t)s.llr_offset + (size_t)s.nof_symbols * qm;
    if ((nof_symbols != 0 && se > nof_symbols) || (llr_bytes != 0 && le > llr_bytes)) {
      return PDC_ERR_INVALID_ARG;
    }
    max_sym = se > max_sym ? se : max_sym, max_llr = le > max_llr ? le : max_llr;
    for (uint32_t first = 0; first < s.nof_symbols; first += 1024) {
      chunks.push_back(i), chunks.push_back(first);
    }
  }
  return PDC_OK;
}

pdc_status pdc_demodulate_device(pdc_ctx* ctx, const pdc_demod_seg* segs, uint32_t n, const float* d_symbols,
                                 const float* d_noise_vars, int8_t* d_llrs, void* stream_v)
{
  if (!ctx || !segs || n == 0 || !d_symbols || !d_noise_vars || !d_llrs) {
    return PDC_ERR_INVALID_ARG;
  }
  std::vector<uint32_t> chunks;
  size_t                max_sym, max_llr;
  pdc_status            st = demod_plan(segs, n, 0, 0, chunks, max_sym, max_llr);
  if (st != PDC_OK) {
    return st;
  }
  if (chunks.empty()) {
    return PDC_OK; // only empty segments: the reference's loops do not run either
  }
  cudaStream_t stream = (cudaStream_t)stream_v;
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> lock(ctx->demod_mtx);
  pdc_ctx::demod_slot&        slot = ctx->demod_ring[ctx->demod_next++ % 8];
  if (!slot.done) {
    CK(cudaEventCreateWithFlags(&slot.done, cudaEventDisableTiming));
  }
  if (slot.used) {
    CK(cudaEventSynchronize(slot.done)); // the kernel that read this staging entry has finished
  }
  const size_t seg_bytes = ((size_t)n * sizeof(pdc_demod_seg) + 15) & ~(size_t)15, chunk_bytes = chunks.size() * 4;
  // Through the slot's pinned image: the caller's array and `chunks` may go away, and the upload does not wait for the stream
  // (a pageable source of more than 64 KB makes cudaMemcpyAsync block until the stream has drained).
  {
    const pdc_status rs = ring_reserve(ctx, ctx->demod_ring, seg_bytes + chunk_bytes, true);
    if (rs != PDC_OK) {
      return rs;
    }
  }
  memcpy(slot.h_desc, segs, (size_t)n * sizeof(pdc_demod_seg));
  memcpy((char*)slot.h_desc + seg_bytes, chunks.data(), chunk_bytes);
  CK(cudaMemcpyAsync(slot.d_desc, slot.h_desc, seg_bytes + chunk_bytes, cudaMemcpyHostToDevice, stream));
  CK(pdc_demod_launch((const pdc_demod_seg*)slot.d_desc, (char*)slot.d_desc + seg_bytes, (uint32_t)(chunks.size() / 2),
                      d_symbols, d_noise_vars, d_llrs, stream));
  ctx->launches += 1;
  CK(cudaEventRecord(slot.done, stream));
  slot.used = true;
  return PDC_OK;
}

pdc_status pdc_demodulate(pdc_ctx* ctx, const pdc_demod_seg* segs, uint32_t n, const float* symbols, const float* noise_vars,
                          size_t nof_symbols, int8_t* llrs, size_t llr_bytes)
{
  if (!ctx || !segs || n == 0 || !symbols || !noise_vars || !llrs || nof_symbols == 0 || llr_bytes == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  std::vector<uint32_t> chunks;
  size_t                max_sym, max_llr;
  pdc_status            st = demod_plan(segs, n, nof_symbols, llr_bytes, chunks, max_sym, max_llr);
  if (st != PDC_OK) {
    return st;
  }
  if (chunks.empty()) {
    return PDC_OK;
  }
  CK(cudaSetDevice(ctx->cfg.device));
  // One call at a time on the context's demodulation stream and buffers (the adapter is synchronous like the reference's
  // demodulate_soft).
  std::lock_guard<std::mutex> call_lock(ctx->demod_call_mtx);
  const size_t                sym4 = (max_sym + 3) & ~(size_t)3; // keeps the noise array 16-byte aligned behind the symbols
  if (!ctx->demod_stream) {
    CK(cudaStreamCreateWithFlags(&ctx->demod_stream, cudaStreamNonBlocking));
  }
  if (ctx->demod_in_cap < sym4 * 3) {
    cudaFree(ctx->d_demod_in);
    ctx->d_demod_in = nullptr, ctx->demod_in_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_demod_in, sym4 * 3 * sizeof(float) * 2));
    ctx->demod_in_cap = sym4 * 3 * 2;
  }
  if (ctx->demod_out_cap < max_llr) {
    cudaFree(ctx->d_demod_out);
    ctx->d_demod_out = nullptr, ctx->demod_out_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_demod_out, max_llr * 2));
    ctx->demod_out_cap = max_llr * 2;
  }
  cudaStream_t stream  = ctx->demod_stream;
  float*       d_in    = ctx->d_demod_in;
  float*       d_noise = d_in + 2 * sym4;
  int8_t*      d_out   = ctx->d_demod_out;
  CK(cudaMemcpyAsync(d_in, symbols, max_sym * 2 * sizeof(float), cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(d_noise, noise_vars, max_sym * sizeof(float), cudaMemcpyHostToDevice, stream));
  st = pdc_demodulate_device(ctx, segs, n, d_in, d_noise, d_out, stream);
  if (st != PDC_OK) {
    return st;
  }
  // Segments may leave gaps in llrs[]: only each segment's own bytes are copied back.
  for (uint32_t i = 0; i != n; ++i) {
    const size_t bytes = (size_t)segs[i].nof_symbols * demod_bits_per_symbol(segs[i].mod);
    if (bytes != 0) {
      CK(cudaMemcpyAsync(llrs + segs[i].llr_offset, d_out + segs[i].llr_offset, bytes, cudaMemcpyDeviceToHost, stream));
    }
  }
  CK(cudaStreamSynchronize(stream));
  return PDC_OK;
}

// ---- UL-SCH demultiplexer --------------------------------------------------------------------------------------
namespace {
// d and the number of REs taken when `remainder` REs are still to place among `avail` candidates
// (ulsch_demultiplex_impl.cpp:361-366 and the four copies of it).
inline void demux_step(uint32_t remainder, uint32_t avail, uint32_t& d, uint32_t& m)
{
  d = 1, m = avail;
  if (remainder < avail) {
    d = avail / remainder, m = remainder;
  }
}

// configure_current_ofdm_symbol (ulsch_demultiplex_impl.cpp:330-463) for every OFDM symbol of the allocation, in closed form:
// the sets are never built on the host, only their sizes. Offsets in `syms` are relative to the codeword / stream starts
// (kept in the pointer fields as integers until the caller rebases them). lens: input, UL-SCH, HARQ-ACK, CSI 1, CSI 2.
pdc_status demux_plan(const pdc_ulsch_demux_desc& c, std::vector<pdc_dev_demux_sym>* syms, uint32_t lens[5], bool& placeholders)
{
  const bool qm_ok = c.Qm == 1 || c.Qm == 2 || c.Qm == 4 || c.Qm == 6 || c.Qm == 8;
  if (!qm_ok || c.nof_layers == 0 || c.nof_layers > 4 || c.nof_prb == 0 || c.nof_prb > 275 || c.nof_symbols == 0 ||
      c.start_symbol_index + c.nof_symbols > 14 || (c.dmrs_type != 1 && c.dmrs_type != 2) || c.nof_cdm_groups_without_data == 0 ||
      c.nof_cdm_groups_without_data > (c.dmrs_type == 1 ? 2 : 3) || (c.dmrs_symbol_mask >> 14) != 0 ||
      (c.flags & ~PDC_DEMUX_F_CSI2_AFTER_CSI1) != 0 || c.reserved != 0 || c.reserved2 != 0) {
    return PDC_ERR_INVALID_ARG;
  }
  // A stream without message bits has no encoded bits; reserved REs only exist for HARQ-ACK of at most two bits (TS 38.212 6.2.7).
  if ((c.nof_harq_ack_bits == 0 && c.nof_enc_harq_ack_bits != 0) || (c.nof_csi_part1_bits == 0 && c.nof_enc_csi_part1_bits != 0) ||
      (c.nof_csi_part2_bits == 0 && c.nof_enc_csi_part2_bits != 0) || (c.nof_harq_ack_bits > 2 && c.nof_harq_ack_rvd != 0)) {
    return PDC_ERR_INVALID_ARG;
  }
  // get_ulsch_demultiplex_l1 / _l1_csi (:32-55): first symbol without DM-RS after the first DM-RS symbol / at all.
  int first_dmrs = -1, l1 = -1, l1_csi = -1;
  for (int l = 0; l != 14; ++l) {
    const bool dm = ((c.dmrs_symbol_mask >> l) & 1) != 0;
    if (dm && first_dmrs < 0) {
      first_dmrs = l;
    }
    if (!dm && l1_csi < 0) {
      l1_csi = l;
    }
    if (!dm && first_dmrs >= 0 && l1 < 0) {
      l1 = l;
    }
  }
  if (first_dmrs < 0 || l1 < 0 || l1_csi < 0) {
    return PDC_ERR_INVALID_ARG; // "No DM-RS symbol found."
  }
  const uint32_t nbpr        = (uint32_t)c.Qm * c.nof_layers;
  const uint32_t re_dmrs_prb = 12u - c.nof_cdm_groups_without_data * (c.dmrs_type == 1 ? 6u : 4u);
  uint32_t       m_rvd = 0, m_harq = 0, m_csi1 = 0, m_csi2 = 0;
  uint32_t       off_in = 0, off_sch = 0;
  bool           csi2_active = c.nof_enc_csi_part2_bits != 0 && (c.flags & PDC_DEMUX_F_CSI2_AFTER_CSI1) == 0;
  placeholders               = false;
  for (uint32_t l = c.start_symbol_index; l != (uint32_t)c.start_symbol_index + c.nof_symbols; ++l) {
    const bool        dm = ((c.dmrs_symbol_mask >> l) & 1) != 0;
    pdc_dev_demux_sym s;
    memset(&s, 0, sizeof(s));
    s.nof_re = (dm ? re_dmrs_prb : 12u) * c.nof_prb, s.nbpr = nbpr, s.Qm = c.Qm;
    s.in = (const int8_t*)(uintptr_t)off_in, s.sch = (int8_t*)(uintptr_t)off_sch;
    s.harq = (int8_t*)(uintptr_t)m_harq, s.csi1 = (int8_t*)(uintptr_t)m_csi1, s.csi2 = (int8_t*)(uintptr_t)m_csi2;
    uint32_t       M_uci   = dm ? 0u : s.nof_re;
    const uint32_t rem_rvd = (c.nof_harq_ack_rvd - m_rvd) / nbpr;
    if ((int)l >= l1 && M_uci > 0 && rem_rvd > 0) { // step 1
      demux_step(rem_rvd, M_uci, s.rvd_d, s.rvd_m);
      m_rvd += s.rvd_m * nbpr;
    }
    const uint32_t rem_harq = (c.nof_enc_harq_ack_bits - m_harq) / nbpr;
    if ((int)l >= l1 && M_uci > 0 && c.nof_harq_ack_bits > 2 && rem_harq > 0) { // step 2
      demux_step(rem_harq, M_uci, s.harq_d, s.harq_m);
      M_uci -= s.harq_m;
      m_harq += s.harq_m * nbpr;
    }
    const uint32_t rem_csi1 = (c.nof_enc_csi_part1_bits - m_csi1) / nbpr;
    const uint32_t M_rvd    = s.rvd_m;
    if ((int)l >= l1_csi && M_uci > M_rvd && rem_csi1 > 0) { // step 3: candidates exclude the reserved REs
      demux_step(rem_csi1, M_uci - M_rvd, s.csi1_d, s.csi1_m);
      M_uci -= s.csi1_m;
      m_csi1 += s.csi1_m * nbpr;
      if (m_csi1 == c.nof_enc_csi_part1_bits && c.nof_enc_csi_part2_bits != 0) {
        csi2_active = true; // the CSI Part 1 feedback configures CSI Part 2 within this symbol (:241-251, :528-533)
      }
    }
    const uint32_t rem_csi2 = (c.nof_enc_csi_part2_bits - m_csi2) / nbpr;
    if (csi2_active && (int)l >= l1_csi && M_uci > 0 && rem_csi2 > 0) { // step 3bis: reserved REs are candidates
      demux_step(rem_csi2, M_uci, s.csi2_d, s.csi2_m);
      m_csi2 += s.csi2_m * nbpr;
    }
    if (M_rvd > 0 && c.nof_harq_ack_bits <= 2 && rem_harq > 0) { // step 5: HARQ-ACK of 1 or 2 bits punctures reserved REs
      demux_step(rem_harq, M_rvd, s.harq_d, s.harq_m);
      s.harq_of_rvd = 1;
      m_harq += s.harq_m * nbpr;
    }
    s.harq_ph = (c.nof_harq_ack_bits == 1 || c.nof_harq_ack_bits == 2) ? (uint8_t)c.nof_harq_ack_bits : 0;
    s.csi1_ph = (c.nof_csi_part1_bits == 1 || c.nof_csi_part1_bits == 2) ? (uint8_t)c.nof_csi_part1_bits : 0;
    s.csi2_ph = (c.nof_csi_part2_bits == 1 || c.nof_csi_part2_bits == 2) ? (uint8_t)c.nof_csi_part2_bits : 0;
    if (c.Qm > 1 && ((s.harq_m && s.harq_ph) || (s.csi1_m && s.csi1_ph) || (s.csi2_m && s.csi2_ph))) {
      placeholders = true;
      if (syms && !pdc_prg_states(c.c_init, off_in, &s.x1_state, &s.x2_state)) {
        return PDC_ERR_INVALID_ARG; // no context has been created yet
      }
    }
    const uint32_t sch_re = s.nof_re - (s.harq_of_rvd ? 0u : s.harq_m) - s.csi1_m - s.csi2_m;
    off_in += s.nof_re * nbpr;
    off_sch += sch_re * nbpr;
    if (syms && s.nof_re != 0) { // a DM-RS symbol without data REs (two CDM groups of type 1) needs no CTA
      syms->push_back(s);
    }
  }
  // on_end_codeword (:321-328) asserts that every UCI stream has been delivered in full.
  if (m_harq != c.nof_enc_harq_ack_bits || m_csi1 != c.nof_enc_csi_part1_bits || m_csi2 != c.nof_enc_csi_part2_bits) {
    return PDC_ERR_INVALID_ARG;
  }
  lens[0] = off_in, lens[1] = off_sch, lens[2] = m_harq, lens[3] = m_csi1, lens[4] = m_csi2;
  return PDC_OK;
}

// Plans every codeword and rebases the symbol records on the buffers. in_len / out_len: host array sizes (0 = unknown).
pdc_status demux_plan_all(const pdc_ulsch_demux_desc* descs, uint32_t n, const int8_t* d_in, int8_t* d_out, size_t in_len,
                          size_t out_len, std::vector<pdc_dev_demux_sym>& syms, bool& placeholders, size_t& max_in, size_t& max_out)
{
  syms.clear();
  placeholders = false;
  max_in = max_out = 0;
  for (uint32_t t = 0; t != n; ++t) {
    const pdc_ulsch_demux_desc& c = descs[t];
    uint32_t                    lens[5];
    bool                        ph    = false;
    const size_t                first = syms.size();
    const pdc_status            st    = demux_plan(c, &syms, lens, ph);
    if (st != PDC_OK) {
      return st;
    }
    placeholders = placeholders || ph;
    const size_t ends[5] = {(size_t)c.in_offset + lens[0], (size_t)c.sch_offset + lens[1], (size_t)c.harq_ack_offset + lens[2],
                            (size_t)c.csi_part1_offset + lens[3], (size_t)c.csi_part2_offset + lens[4]};
    max_in               = std::max(max_in, ends[0]);
    for (int k = 1; k != 5; ++k) {
      if (lens[k] != 0) {
        max_out = std::max(max_out, ends[k]);
      }
    }
    if ((in_len != 0 && max_in > in_len) || (out_len != 0 && max_out > out_len)) {
      return PDC_ERR_INVALID_ARG;
    }
    for (size_t q = first; q != syms.size(); ++q) {
      pdc_dev_demux_sym& s = syms[q];
      s.in   = d_in + c.in_offset + (uintptr_t)s.in;
      s.sch  = d_out + c.sch_offset + (uintptr_t)s.sch;
      s.harq = d_out + c.harq_ack_offset + (uintptr_t)s.harq;
      s.csi1 = d_out + c.csi_part1_offset + (uintptr_t)s.csi1;
      s.csi2 = d_out + c.csi_part2_offset + (uintptr_t)s.csi2;
    }
  }
  return PDC_OK;
}
} // namespace

extern "C" {

pdc_status pdc_ulsch_demux_lengths(const pdc_ulsch_demux_desc* d, uint32_t lens[5])
{
  if (!d || !lens) {
    return PDC_ERR_INVALID_ARG;
  }
  bool ph = false;
  return demux_plan(*d, nullptr, lens, ph);
}

pdc_status pdc_ulsch_demultiplex_device(pdc_ctx* ctx, const pdc_ulsch_demux_desc* descs, uint32_t n, const int8_t* d_in,
                                        int8_t* d_out, void* stream_v)
{
  if (!ctx || !descs || n == 0 || !d_in || !d_out) {
    return PDC_ERR_INVALID_ARG;
  }
  std::vector<pdc_dev_demux_sym> syms;
  bool                           placeholders;
  size_t                         max_in, max_out;
  const pdc_status               st = demux_plan_all(descs, n, d_in, d_out, 0, 0, syms, placeholders, max_in, max_out);
  if (st != PDC_OK) {
    return st;
  }
  // Symbols without UCI first: they go to the copy kernel, the others to the kernel that builds the RE sets.
  const size_t nplain = std::stable_partition(syms.begin(), syms.end(), [](const pdc_dev_demux_sym& s) {
                          return s.harq_m == 0 && s.csi1_m == 0 && s.csi2_m == 0;
                        }) - syms.begin();
  cudaStream_t stream = (cudaStream_t)stream_v;
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> lock(ctx->demod_mtx);
  pdc_ctx::demod_slot&        slot = ctx->demux_ring[ctx->demux_next++ % 8];
  if (!slot.done) {
    CK(cudaEventCreateWithFlags(&slot.done, cudaEventDisableTiming));
  }
  if (slot.used) {
    CK(cudaEventSynchronize(slot.done)); // the kernel that read this staging entry has finished
  }
  const size_t bytes = syms.size() * sizeof(pdc_dev_demux_sym);
  {
    const pdc_status rs = ring_reserve(ctx, ctx->demux_ring, bytes, true);
    if (rs != PDC_OK) {
      return rs;
    }
  }
  memcpy(slot.h_desc, syms.data(), bytes); // pinned image: the upload does not wait for the stream to drain
  CK(cudaMemcpyAsync(slot.d_desc, slot.h_desc, bytes, cudaMemcpyHostToDevice, stream));
  CK(pdc_demux_launch((const pdc_dev_demux_sym*)slot.d_desc, (uint32_t)nplain, (uint32_t)(syms.size() - nplain), placeholders, stream));
  ctx->launches += (nplain != 0 ? 1 : 0) + (nplain != syms.size() ? 1 : 0);
  CK(cudaEventRecord(slot.done, stream));
  slot.used = true;
  return PDC_OK;
}

pdc_status pdc_ulsch_demultiplex(pdc_ctx* ctx, const pdc_ulsch_demux_desc* descs, uint32_t n, const int8_t* in, size_t in_len,
                                 int8_t* out, size_t out_len)
{
  if (!ctx || !descs || n == 0 || !in || !out || in_len == 0 || out_len == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  std::vector<pdc_dev_demux_sym> syms;
  bool                           placeholders;
  size_t                         max_in, max_out;
  pdc_status                     st = demux_plan_all(descs, n, nullptr, nullptr, in_len, out_len, syms, placeholders, max_in, max_out);
  if (st != PDC_OK) {
    return st;
  }
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> call_lock(ctx->demod_call_mtx); // one host-buffer call at a time on the context's stream
  if (!ctx->demod_stream) {
    CK(cudaStreamCreateWithFlags(&ctx->demod_stream, cudaStreamNonBlocking));
  }
  if (ctx->demux_in_cap < max_in) {
    cudaFree(ctx->d_demux_in);
    ctx->d_demux_in = nullptr, ctx->demux_in_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_demux_in, max_in * 2));
    ctx->demux_in_cap = max_in * 2;
  }
  if (ctx->demux_out_cap < max_out + 4) {
    cudaFree(ctx->d_demux_out);
    ctx->d_demux_out = nullptr, ctx->demux_out_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_demux_out, (max_out + 4) * 2));
    ctx->demux_out_cap = (max_out + 4) * 2;
  }
  cudaStream_t stream = ctx->demod_stream;
  CK(cudaMemcpyAsync(ctx->d_demux_in, in, max_in, cudaMemcpyHostToDevice, stream));
  st = pdc_ulsch_demultiplex_device(ctx, descs, n, ctx->d_demux_in, ctx->d_demux_out, stream);
  if (st != PDC_OK) {
    return st;
  }
  if (max_out != 0) {
    // The streams tile disjoint ranges of `out`; bytes between them are not the library's to touch: copy stream by stream.
    for (uint32_t t = 0; t != n; ++t) {
      uint32_t lens[5];
      bool     ph = false;
      demux_plan(descs[t], nullptr, lens, ph);
      const uint64_t offs[4] = {descs[t].sch_offset, descs[t].harq_ack_offset, descs[t].csi_part1_offset, descs[t].csi_part2_offset};
      for (int k = 0; k != 4; ++k) {
        if (lens[k + 1] != 0) {
          CK(cudaMemcpyAsync(out + offs[k], ctx->d_demux_out + offs[k], lens[k + 1], cudaMemcpyDeviceToHost, stream));
        }
      }
    }
  }
  CK(cudaStreamSynchronize(stream));
  return PDC_OK;
}

} // extern "C"

// ---- PDSCH transport-block encoder ----------------------------------------------------------------------------
namespace {
struct enc_seg {
  int bg, Z, C, K, Kp, F, tb_crc_bits, cb_crc_bits;
};
// TS 38.212 5.2.2 as the reference derives it (ldpc.h:128-214, ldpc_segmenter_tx_impl.cpp): number of code blocks, lifting
// size, segment length K' (every code block carries ceil(B' / C) bits, the transport block is zero-padded), fillers.
bool enc_segmentation(int bg, uint32_t tbs, enc_seg& s)
{
  static const int sizes[PDC_NOF_LIFTING_SIZES] = {2,   3,   4,   5,   6,   7,   8,   9,   10,  11,  12,  13,  14,
                                                   15,  16,  18,  20,  22,  24,  26,  28,  30,  32,  36,  40,  44,
                                                   48,  52,  56,  60,  64,  72,  80,  88,  96,  104, 112, 120, 128,
                                                   144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};
  s.bg          = bg;
  s.tb_crc_bits = (tbs > 3824) ? 24 : 16;
  const int B   = (int)tbs + s.tb_crc_bits;
  const int Kcb = (bg == 1) ? 8448 : 3840;
  if (B <= Kcb) {
    s.C = 1, s.cb_crc_bits = 0;
  } else {
    s.C           = (B + (Kcb - 24) - 1) / (Kcb - 24);
    s.cb_crc_bits = 24;
  }
  const int Bp = B + s.C * s.cb_crc_bits;
  s.Kp         = (Bp + s.C - 1) / s.C;
  int Kb       = 22;
  if (bg == 2) {
    Kb = (B > 640) ? 10 : ((B > 560) ? 9 : ((B > 192) ? 8 : 6));
  }
  s.Z = -1;
  for (int i = 0; i != PDC_NOF_LIFTING_SIZES; ++i) {
    if (Kb * sizes[i] >= s.Kp) {
      s.Z = sizes[i];
      break;
    }
  }
  if (s.Z < 0) {
    return false;
  }
  s.K = ((bg == 1) ? 22 : 10) * s.Z;
  s.F = s.K - s.Kp;
  return true;
}
} // namespace

static pdc_status enc_plan(pdc_ctx* ctx, const pdc_pdsch_tb_desc* tbs, uint32_t n, size_t tb_len, size_t cw_len,
                           const uint8_t* d_tb, uint8_t* d_cw, uint32_t* d_crc_acc, std::vector<pdc_dev_enc_tb>& out,
                           size_t& smem, size_t& max_tb, size_t& max_cw, uint32_t& max_cbs)
{
  out.clear();
  smem = max_tb = max_cw = 0;
  max_cbs = 0;
  for (uint32_t t = 0; t != n; ++t) {
    const pdc_pdsch_tb_desc& d = tbs[t];
    enc_seg                  s;
    const bool qm_ok = d.Qm == 1 || d.Qm == 2 || d.Qm == 4 || d.Qm == 6 || d.Qm == 8;
    if ((d.bg != 1 && d.bg != 2) || !qm_ok || d.rv > 3 || d.tbs_bits == 0 || (d.tbs_bits & 7) != 0 || d.tbs_bits + 24u > 1277992u ||
        d.reserved != 0 || !enc_segmentation(d.bg, d.tbs_bits, s)) {
      return PDC_ERR_INVALID_ARG; // ldpc_segmenter_tx_impl.cpp:40-55
    }
    // Rate-matched lengths: derived from the channel symbols (ldpc_segmenter_helpers.h:82-100: the first C - rem code blocks
    // get floor, the others ceil), or taken as given (nof_ch_symbols == 0: hal::hw_pdsch_encoder_configuration carries
    // cw_length_a / cw_length_b / nof_short_segments and no layer count, hw_accelerator_pdsch_enc.h:44-51).
    uint32_t nof_short = 0, E_short = 0, E_long = 0;
    if (d.nof_ch_symbols != 0) {
      if (d.nof_layers == 0 || d.nof_layers > 4 || (d.nof_ch_symbols % d.nof_layers) != 0 || d.cw_length_a != 0 || d.cw_length_b != 0 ||
          d.nof_short_segments != 0 || (uint64_t)(d.nof_ch_symbols / d.nof_layers) < (uint64_t)s.C) {
        return PDC_ERR_INVALID_ARG; // every code block needs at least one channel symbol
      }
      const uint32_t sym = d.nof_ch_symbols / d.nof_layers, base = sym / s.C, rem = sym % s.C;
      nof_short = (uint32_t)s.C - rem;
      E_short = (uint32_t)d.nof_layers * d.Qm * base, E_long = (uint32_t)d.nof_layers * d.Qm * (base + 1);
    } else {
      nof_short = d.nof_short_segments, E_short = d.cw_length_a, E_long = d.cw_length_b;
      if (nof_short == 0 || nof_short > (uint32_t)s.C || E_short == 0 || (E_short % d.Qm) != 0 || (E_long % d.Qm) != 0 ||
          (nof_short < (uint32_t)s.C && E_long == 0)) {
        return PDC_ERR_INVALID_ARG;
      }
      E_long = (nof_short == (uint32_t)s.C) ? E_short : E_long;
    }
    const uint64_t G = (uint64_t)nof_short * E_short + (uint64_t)((uint32_t)s.C - nof_short) * E_long; // == nof_ch_symbols * Qm
    const size_t   te = (size_t)d.tb_offset + d.tbs_bits / 8, ce = (size_t)d.cw_offset + (size_t)G;
    if ((tb_len != 0 && te > tb_len) || (cw_len != 0 && ce > cw_len)) {
      return PDC_ERR_INVALID_ARG;
    }
    max_tb = te > max_tb ? te : max_tb, max_cw = ce > max_cw ? ce : max_cw;
    const int Kb = (d.bg == 1) ? 22 : 10, Nfull = (d.bg == 1) ? 68 : 52, M = Nfull - Kb;
    const int N = (Nfull - 2) * s.Z, Ncb = (d.Nref > 0) ? ((int)d.Nref < N ? (int)d.Nref : N) : N;
    const int n_sys = (Kb - 2) * s.Z, n_info = n_sys - s.F;
    const int fw = std::max(0, std::min(n_sys, Ncb) - std::min(n_info, Ncb)), Lc = Ncb - fw;
    static const int f1[4] = {0, 17, 33, 56}, f2[4] = {0, 13, 25, 43};
    const int fr = (d.bg == 1) ? f1[d.rv] : f2[d.rv];
    const int k0 = (int)(((long long)fr * Ncb) / N) * s.Z;
    const int t0 = (k0 < n_info) ? k0 : ((k0 >= n_sys) ? k0 - fw : n_info);
    if (Lc <= 0) {
      return PDC_ERR_INVALID_ARG;
    }
    // Column Kb in the four core rows: three circulants, two of them equal (TS 38.212 tables 5.3.2-2/3); the odd one out is the
    // shift b with which the sum of the four rows leaves P^b p1 (ldpc_encoder_impl.cpp, high-rate region).
    int core_shift[4] = {-1, -1, -1, -1}, bshift = 0;
    {
      const uint8_t*  deg = (d.bg == 1) ? pdc_bg1_row_deg : pdc_bg2_row_deg;
      const uint8_t*  col = (d.bg == 1) ? pdc_bg1_edge_col : pdc_bg2_edge_col;
      const uint16_t* sh  = (d.bg == 1) ? pdc_bg1_edge_shift[lifting_set_index(s.Z)] : pdc_bg2_edge_shift[lifting_set_index(s.Z)];
      int             e   = 0;
      for (int m = 0; m != 4; ++m) {
        for (int k = 0; k != deg[m]; ++k, ++e) {
          if (col[e] == Kb) {
            core_shift[m] = sh[e] % s.Z;
          }
        }
      }
      for (int m = 0; m != 4; ++m) {
        int cnt = 0;
        for (int k = 0; k != 4; ++k) {
          cnt += (core_shift[m] >= 0 && core_shift[k] == core_shift[m]);
        }
        if (cnt & 1) {
          bshift = core_shift[m];
        }
      }
    }
    auto rows_for = [&](uint32_t E) {
      if ((long long)t0 + E <= Lc) {
        const int tmax = t0 + (int)E - 1, idx = (tmax < n_info) ? tmax : tmax + fw;
        return std::max(4, std::min(M, (idx / s.Z + 1) + 2 - Kb));
      }
      return std::max(4, std::min(M, (Ncb + s.Z - 1) / s.Z + 2 - Kb));
    };
    pdc_dev_enc_tb c;
    memset(&c, 0, sizeof(c));
    c.tb = d_tb + d.tb_offset, c.out = d_cw + d.cw_offset, c.tb_crc = d_crc_acc + t;
    c.tbs_bits = d.tbs_bits, c.nof_bits = (uint32_t)(s.Kp - s.cb_crc_bits);
    c.nof_cbs = (uint32_t)s.C, c.nof_short = nof_short;
    c.E_short = E_short, c.E_long = E_long;
    c.Nref = d.Nref, c.nof_filler = (uint32_t)s.F;
    c.Z = (uint16_t)s.Z, c.graph_idx = (uint16_t)((d.bg - 1) * PDC_NOF_LIFTING_SIZES + pdc_lifting_pos(s.Z));
    c.bg = d.bg, c.Qm = d.Qm, c.rv = d.rv, c.tb_crc_type = (s.tb_crc_bits == 24) ? PDC_CRC24A : PDC_CRC16;
    c.cb_crc = s.cb_crc_bits ? 1 : 0;
    c.rows_short = (uint8_t)rows_for(c.E_short), c.rows_long = (uint8_t)rows_for(c.E_long);
    c.bshift = (uint16_t)bshift;
    for (int m = 0; m != 4; ++m) {
      c.core_shift[m] = (int16_t)core_shift[m];
    }
    smem    = std::max(smem, (size_t)(Kb + std::max(c.rows_short, c.rows_long) + 5) * s.Z);
    max_cbs = std::max(max_cbs, (uint32_t)s.C);
    out.push_back(c);
  }
  (void)ctx;
  return PDC_OK;
}

pdc_status pdc_pdsch_encode_device(pdc_ctx* ctx, const pdc_pdsch_tb_desc* tbs, uint32_t n, const uint8_t* d_tb_bytes,
                                   uint8_t* d_codewords, void* stream_v)
{
  if (!ctx || !tbs || n == 0 || !d_tb_bytes || !d_codewords) {
    return PDC_ERR_INVALID_ARG;
  }
  cudaStream_t stream = (cudaStream_t)stream_v;
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> lock(ctx->demod_mtx);
  pdc_ctx::demod_slot&        slot = ctx->enc_ring[ctx->enc_next++ % 8];
  if (!slot.done) {
    CK(cudaEventCreateWithFlags(&slot.done, cudaEventDisableTiming));
  }
  if (slot.used) {
    CK(cudaEventSynchronize(slot.done));
  }
  // Device staging: n CRC accumulators, then one descriptor per transport block (the kernels derive their code block).
  std::vector<pdc_dev_enc_tb> cbs;
  size_t                      smem, max_tb, max_cw;
  uint32_t                    max_cbs;
  const size_t                acc_bytes = ((size_t)n * 4 + 15) & ~(size_t)15;
  const size_t                need      = acc_bytes + (size_t)n * sizeof(pdc_dev_enc_tb);
  {
    const pdc_status rs = ring_reserve(ctx, ctx->enc_ring, need, false);
    if (rs != PDC_OK) {
      return rs;
    }
  }
  uint32_t*  d_acc = (uint32_t*)slot.d_desc;
  pdc_status st    = enc_plan(ctx, tbs, n, 0, 0, d_tb_bytes, d_codewords, d_acc, cbs, smem, max_tb, max_cw, max_cbs);
  if (st != PDC_OK) {
    return st;
  }
  pdc_dev_enc_tb* d_tbs = (pdc_dev_enc_tb*)((char*)slot.d_desc + acc_bytes);
  CK(cudaMemsetAsync(d_acc, 0, acc_bytes, stream));
  CK(cudaMemcpyAsync(d_tbs, cbs.data(), cbs.size() * sizeof(pdc_dev_enc_tb), cudaMemcpyHostToDevice, stream));
  CK(pdc_enc_launch(d_tbs, n, max_cbs, ctx->d_graphs, ctx->d_crc, smem, stream));
  ctx->launches += 2;
  CK(cudaEventRecord(slot.done, stream));
  slot.used = true;
  return PDC_OK;
}

pdc_status pdc_pdsch_encode(pdc_ctx* ctx, const pdc_pdsch_tb_desc* tbs, uint32_t n, const uint8_t* tb_bytes, size_t tb_len,
                            uint8_t* codewords, size_t cw_len)
{
  if (!ctx || !tbs || n == 0 || !tb_bytes || !codewords || tb_len == 0 || cw_len == 0) {
    return PDC_ERR_INVALID_ARG;
  }
  // Validate against the host array sizes first (no device pointers needed for that).
  std::vector<pdc_dev_enc_tb> cbs;
  size_t                      smem, max_tb, max_cw;
  uint32_t                    max_cbs;
  pdc_status st = enc_plan(ctx, tbs, n, tb_len, cw_len, nullptr, nullptr, nullptr, cbs, smem, max_tb, max_cw, max_cbs);
  if (st != PDC_OK) {
    return st;
  }
  CK(cudaSetDevice(ctx->cfg.device));
  std::lock_guard<std::mutex> call_lock(ctx->demod_call_mtx);
  if (!ctx->demod_stream) {
    CK(cudaStreamCreateWithFlags(&ctx->demod_stream, cudaStreamNonBlocking));
  }
  if (ctx->enc_in_cap < max_tb) {
    cudaFree(ctx->d_enc_in);
    ctx->d_enc_in = nullptr, ctx->enc_in_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_enc_in, max_tb * 2));
    ctx->enc_in_cap = max_tb * 2;
  }
  if (ctx->enc_out_cap < max_cw) {
    cudaFree(ctx->d_enc_out);
    ctx->d_enc_out = nullptr, ctx->enc_out_cap = 0;
    CK(cudaMalloc((void**)&ctx->d_enc_out, max_cw * 2));
    ctx->enc_out_cap = max_cw * 2;
  }
  cudaStream_t stream = ctx->demod_stream;
  CK(cudaMemcpyAsync(ctx->d_enc_in, tb_bytes, max_tb, cudaMemcpyHostToDevice, stream));
  st = pdc_pdsch_encode_device(ctx, tbs, n, ctx->d_enc_in, ctx->d_enc_out, stream);
  if (st != PDC_OK) {
    return st;
  }
  for (uint32_t t = 0; t != n; ++t) {
    const size_t G = (size_t)cbs[t].nof_short * cbs[t].E_short + (size_t)(cbs[t].nof_cbs - cbs[t].nof_short) * cbs[t].E_long;
    CK(cudaMemcpyAsync(codewords + tbs[t].cw_offset, ctx->d_enc_out + tbs[t].cw_offset, G, cudaMemcpyDeviceToHost, stream));
  }
  CK(cudaStreamSynchronize(stream));
  return PDC_OK;
}

pdc_status pdc_harq_free(pdc_ctx* ctx, int32_t slot)
{
  if (!ctx || slot < 0 || (uint32_t)slot >= ctx->cfg.harq_slots) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaMemset(ctx->d_harq + (size_t)slot * PDC_MAX_N, 0, PDC_MAX_N));
  ctx->harq_nz[slot]       = 0;
  ctx->harq_released[slot] = 0;
  return PDC_OK;
}

pdc_status pdc_harq_release(pdc_ctx* ctx, int32_t slot)
{
  if (!ctx || slot < 0 || (uint32_t)slot >= ctx->cfg.harq_slots) {
    return PDC_ERR_INVALID_ARG;
  }
  // Host-side only: the next operation on the slot either rewrites the whole buffer (a first transmission) or has the slot
  // cleared on its stream first (plan_batch / clear_released_slots).
  if (ctx->harq_nz[slot] != 0) {
    ctx->harq_released[slot] = 1;
  }
  return PDC_OK;
}

pdc_status pdc_harq_read(pdc_ctx* ctx, int32_t slot, int8_t* host_dst, size_t n)
{
  if (!ctx || !host_dst || slot < 0 || (uint32_t)slot >= ctx->cfg.harq_slots || n > PDC_MAX_N) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  if (ctx->harq_released[slot]) { // reads as zeros: clear it for real now
    CK(cudaMemset(ctx->d_harq + (size_t)slot * PDC_MAX_N, 0, PDC_MAX_N));
    ctx->harq_nz[slot] = 0, ctx->harq_released[slot] = 0;
  }
  CK(cudaMemcpy(host_dst, ctx->d_harq + (size_t)slot * PDC_MAX_N, n, cudaMemcpyDeviceToHost));
  return PDC_OK;
}

pdc_status pdc_harq_write(pdc_ctx* ctx, int32_t slot, const int8_t* host_src, size_t n)
{
  if (!ctx || !host_src || slot < 0 || (uint32_t)slot >= ctx->cfg.harq_slots || n > PDC_MAX_N) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  if (ctx->harq_released[slot]) {
    CK(cudaMemset(ctx->d_harq + (size_t)slot * PDC_MAX_N, 0, PDC_MAX_N));
    ctx->harq_nz[slot] = 0, ctx->harq_released[slot] = 0;
  }
  CK(cudaMemcpy(ctx->d_harq + (size_t)slot * PDC_MAX_N, host_src, n, cudaMemcpyHostToDevice));
  ctx->harq_nz[slot] = std::max(ctx->harq_nz[slot], (int32_t)n);
  return PDC_OK;
}

uint64_t pdc_kernel_launches(const pdc_ctx* ctx)
{
  return ctx ? ctx->launches.load() : 0;
}

pdc_status pdc_probe_issue_rate(pdc_ctx* ctx, int mode, double* thread_inst_per_s)
{
  if (!ctx || !thread_inst_per_s) {
    return PDC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(ctx->mtx);
  CK(cudaSetDevice(ctx->cfg.device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, ctx->cfg.device));
  CK(pdc_probe_run(mode, prop.multiProcessorCount, ctx->tickets[0].stream, thread_inst_per_s));
  ctx->launches += 4;
  return PDC_OK;
}

uint32_t pdc_cb_message_bits(int bg, int Z)
{
  pdc_geom g;
  return (pdc_geom_init(g, bg, Z) && pdc_lifting_pos(Z) >= 0) ? (uint32_t)g.K : 0;
}

uint32_t pdc_cb_buffer_len(int bg, int Z)
{
  pdc_geom g;
  return (pdc_geom_init(g, bg, Z) && pdc_lifting_pos(Z) >= 0) ? (uint32_t)g.N : 0;
}

uint32_t pdc_cb_max_layers(const pdc_cb_desc* d)
{
  pdc_geom g;
  int      ml = 0;
  if (!d || !validate_desc(*d, g, ml)) {
    return 0;
  }
  return (uint32_t)ml;
}

pdc_status pdc_pdsch_segmentation(int bg, uint32_t tbs_bits, uint32_t* nof_cbs, uint32_t* lifting_size, uint32_t* cb_info_bits,
                                  uint32_t* nof_filler_bits)
{
  enc_seg s;
  if ((bg != 1 && bg != 2) || tbs_bits == 0 || (tbs_bits & 7) != 0 || tbs_bits + 24u > 1277992u || !enc_segmentation(bg, tbs_bits, s)) {
    return PDC_ERR_INVALID_ARG;
  }
  if (nof_cbs) {
    *nof_cbs = (uint32_t)s.C;
  }
  if (lifting_size) {
    *lifting_size = (uint32_t)s.Z;
  }
  if (cb_info_bits) {
    *cb_info_bits = (uint32_t)s.Kp;
  }
  if (nof_filler_bits) {
    *nof_filler_bits = (uint32_t)s.F;
  }
  return PDC_OK;
}

} // extern "C"
