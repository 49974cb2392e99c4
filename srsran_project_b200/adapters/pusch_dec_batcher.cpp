// See pusch_dec_batcher.h.
#include "pusch_dec_batcher.h"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <thread>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

using namespace srsran::cuda;
using clk = std::chrono::steady_clock;

// Copy into the page-locked region with streaming stores: the destination is read next by the DMA engine, not by this core, and
// 1.4 MB per transport block and thread through the caches would evict the working set of the host code around the
// accelerator (the reference's segmenter, CRC checks and assembly are memory-bound already). dst is 64-byte aligned.
static void copy_streaming(int8_t* dst, const int8_t* src, size_t n)
{
#if defined(__SSE2__)
  size_t i = 0;
  if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    for (; i + 64 <= n; i += 64) {
      const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
      const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 16));
      const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 32));
      const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 48));
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), a);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 16), b);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 32), c);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 48), d);
    }
  }
  std::memcpy(dst + i, src + i, n - i);
#else
  std::memcpy(dst, src, n);
#endif
}

struct pusch_dec_batcher::batch {
  enum class st { idle, busy, done };
  st                         state = st::idle;
  std::vector<part*>         members;
  unsigned                   released = 0;
  std::vector<pdc_cb_desc>   descs;
  std::vector<pdc_cb_result> results;
  std::vector<uint8_t>       bits;
  std::vector<int8_t>        soft;
  pdc_ticket                 ticket = 0;
};

pusch_dec_batcher::pusch_dec_batcher(std::shared_ptr<pusch_dec_context> context, const pusch_dec_batcher_config& cfg_) :
  ctx(std::move(context)), cfg(cfg_)
{
  if (!ctx) {
    throw std::runtime_error("pusch_dec_batcher: an accelerator context is required");
  }
  const pusch_dec_context_config& c = ctx->config();
  batch_cbs                         = c.max_batch_cbs;
  cfg.max_parts                     = std::max(1u, cfg.max_parts);
  cfg.nof_batches                   = std::max(1u, std::min(cfg.nof_batches, c.nof_instances));
  cfg.eager_inflight = std::max(1u, cfg.eager_inflight);
  // One region per part; a region holds what one part may bring: the context's batch capacity shared out over max_parts, but
  // never less than the largest rate-matched code block (35 x 8448 soft bits: ldpc.h MAX_CODEBLOCK_RM_SIZE).
  part_bytes = std::max<size_t>(c.max_batch_llrs / cfg.max_parts, 35u * 8448u + 64u);
  part_bytes = (part_bytes + 255) & ~(size_t)255;
  // Parts being filled + ready + in flight: two generations per batch buffer.
  const unsigned nof_parts = 2 * cfg.max_parts * cfg.nof_batches;
  if (pdc_stage_open(ctx->get(), part_bytes * nof_parts, &stage) != PDC_OK) {
    throw std::runtime_error(std::string("pusch_dec_batcher: staging arena could not be allocated: ") + pdc_last_error(ctx->get()));
  }
  parts.resize(nof_parts);
  int8_t* base = pdc_stage_host(stage);
  for (unsigned i = 0; i != nof_parts; ++i) {
    parts[i].llrs     = base + (size_t)i * part_bytes;
    parts[i].llr_base = (size_t)i * part_bytes;
  }
  for (unsigned i = 0; i != cfg.nof_batches; ++i) {
    auto b = std::make_unique<batch>();
    b->descs.reserve(batch_cbs);
    b->results.resize(batch_cbs);
    b->bits.resize((size_t)batch_cbs * BITS_STRIDE);
    batches.push_back(std::move(b));
  }
}

pusch_dec_batcher::~pusch_dec_batcher()
{
  // Every user holds a shared_ptr to the batcher: nothing is in flight any more when the last one lets go.
  pdc_stage_close(stage);
}

pusch_dec_batcher::part* pusch_dec_batcher::join(unsigned max_cbs, size_t max_llr_bytes, unsigned wait_us)
{
  max_cbs       = std::max(1u, std::min(max_cbs, batch_cbs));
  max_llr_bytes = std::min((max_llr_bytes + 63) & ~(size_t)63, part_bytes);
  std::unique_lock<std::mutex> lock(mtx);
  for (int attempt = 0; attempt != 2; ++attempt) {
    part* p = join_locked(max_cbs, max_llr_bytes);
    if (p || wait_us == 0 || attempt == 1) {
      return p;
    }
    const clk::time_point t0 = clk::now();
    cv_room.wait_for(lock, std::chrono::microseconds(wait_us));
    tm.join_wait_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t0).count();
  }
  return nullptr;
}

pusch_dec_batcher::part* pusch_dec_batcher::join_locked(unsigned max_cbs, size_t max_llr_bytes)
{
  for (part& p : parts) {
    if (p.state == part::st::idle) {
      p.state    = part::st::filling;
      p.owner    = nullptr;
      p.llr_cap  = max_llr_bytes;
      p.llr_used = 0;
      p.llr_pushed = 0;
      p.cb_cap   = max_cbs;
      p.descs.clear();
      p.soft.clear();
      p.index.clear();
      p.seen_done = false;
      return &p;
    }
  }
  return nullptr;
}

bool pusch_dec_batcher::add(part* p, pdc_cb_desc d, const int8_t* llrs, size_t nof_llrs, const int8_t* soft_in, size_t nof_soft)
{
  // Only the owning thread touches a part that is being filled: no lock.
  const size_t off = (p->llr_used + 63) & ~(size_t)63; // whole cache lines per code block (streaming stores)
  if (p->state != part::st::filling || p->descs.size() >= p->cb_cap || off + nof_llrs > p->llr_cap) {
    return false;
  }
  // A descriptor the runtime would reject (the reference's srsran_assert contracts) must not take the other callers' code
  // blocks down with it: it is answered with status 1 and never submitted.
  d.llr_offset = d.out_offset = d.soft_offset = 0;
  if (pdc_cb_max_layers(&d) == 0) {
    p->descs.push_back(d);
    p->index.push_back(-1);
    return true;
  }
  const clk::time_point t0 = clk::now();
  copy_streaming(p->llrs + off, llrs, nof_llrs);
  add_ns.fetch_add(std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t0).count(), std::memory_order_relaxed);
  p->llr_used  = off + nof_llrs;
  d.llr_offset = p->llr_base + off;
  if (cfg.push_chunk_bytes != 0 && p->llr_used - p->llr_pushed >= cfg.push_chunk_bytes) {
    push_pending(p);
  }
  if (d.harq_slot < 0 && !(d.flags & PDC_F_DECODE_ONLY)) {
    const size_t so = p->soft.size();
    p->soft.resize(so + SOFT_STRIDE, 0);
    if (soft_in && nof_soft) {
      std::memcpy(p->soft.data() + so, soft_in, std::min(nof_soft, SOFT_STRIDE));
    }
  }
  p->descs.push_back(d);
  p->index.push_back(0);
  return true;
}

void pusch_dec_batcher::push_pending(part* p)
{
  if (p->llr_used == p->llr_pushed) {
    return;
  }
#if defined(__SSE2__)
  _mm_sfence(); // the streaming stores of add() are globally visible before the DMA is told to read them
#endif
  pdc_stage_push(stage, p->llr_base + p->llr_pushed, p->llr_used - p->llr_pushed);
  p->llr_pushed = p->llr_used;
}

void pusch_dec_batcher::seal(part* p)
{
  if (p->state != part::st::filling) {
    return;
  }
  // The rest of the region travels to the device now, while the other members of its future batch are still being filled.
  push_pending(p);
  std::lock_guard<std::mutex> lock(mtx);
  p->state     = part::st::ready;
  p->sealed_at = clk::now();
  ready.push_back(p);
}

void pusch_dec_batcher::fail_batch(batch* b)
{
  for (size_t k = 0; k != b->descs.size(); ++k) {
    b->results[k].iterations = -1, b->results[k].crc_pass = 0, b->results[k].status = 1, b->results[k].nof_layers = 0;
  }
}

// Forms a batch from the ready parts if the closing condition holds, runs it and waits for it (lock released meanwhile).
// Returns true if a batch was run.
bool pusch_dec_batcher::try_form(std::unique_lock<std::mutex>& lock)
{
  if (ready.empty()) {
    return false;
  }
  const bool timeout = clk::now() - ready.front()->sealed_at >= std::chrono::microseconds(cfg.close_after_us);
  if (!(ready.size() >= cfg.max_parts || inflight < cfg.eager_inflight || timeout)) {
    return false;
  }
  batch* b = nullptr;
  for (auto& cand : batches) {
    if (cand->state == batch::st::idle) {
      b = cand.get();
      break;
    }
  }
  if (!b) {
    return false; // every result buffer is still being read by its parts' owners
  }
  // Oldest parts first, as many as fit.
  b->state = batch::st::busy;
  b->members.clear();
  b->released = 0;
  b->descs.clear();
  size_t nsoft = 0;
  size_t taken = 0;
  const clk::time_point formed_at = clk::now();
  while (taken != ready.size() && b->members.size() != cfg.max_parts) {
    part* p = ready[taken];
    if (b->descs.size() + p->descs.size() > batch_cbs && !b->members.empty()) {
      break;
    }
    size_t ps = 0;
    for (size_t k = 0; k != p->descs.size(); ++k) {
      if (p->index[k] < 0) {
        continue;
      }
      pdc_cb_desc d = p->descs[k];
      p->index[k]   = (int)b->descs.size();
      d.out_offset  = b->descs.size() * BITS_STRIDE;
      if (d.harq_slot < 0 && !(d.flags & PDC_F_DECODE_ONLY)) {
        d.soft_offset = nsoft * SOFT_STRIDE;
        if (b->soft.size() < (nsoft + 1) * SOFT_STRIDE) {
          b->soft.resize((nsoft + 1) * SOFT_STRIDE);
        }
        std::memcpy(b->soft.data() + d.soft_offset, p->soft.data() + ps, SOFT_STRIDE);
        ps += SOFT_STRIDE;
        ++nsoft;
      }
      b->descs.push_back(d);
    }
    p->state = part::st::inflight;
    p->owner = b;
    tm.queue_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(formed_at - p->sealed_at).count();
    ++tm.parts;
    b->members.push_back(p);
    ++taken;
  }
  ready.erase(ready.begin(), ready.begin() + taken);
  const unsigned n = (unsigned)b->descs.size();
  ++inflight;
  stat_batches++;
  stat_cbs += n;
  lock.unlock();

  pdc_status st = PDC_OK;
  clk::time_point t_sub = clk::now(), t_mid = t_sub;
  if (n != 0) {
    while ((st = pdc_submit_staged(ctx->get(), stage, b->descs.data(), n, nsoft ? b->soft.data() : nullptr, nsoft * SOFT_STRIDE,
                                   &b->ticket)) == PDC_ERR_CAPACITY) {
      std::this_thread::yield(); // every ticket of the context is in flight (other users of the context)
    }
    t_mid = clk::now();
    if (st == PDC_OK) {
      st = pdc_fetch(ctx->get(), b->ticket, b->bits.data(), b->bits.size(), b->results.data(), nsoft ? b->soft.data() : nullptr,
                     b->soft.size());
    }
    if (st != PDC_OK) {
      std::fprintf(stderr, "pusch_dec_batcher: batch of %u code blocks failed (status %d: %s)\n", n, (int)st,
                   pdc_last_error(ctx->get()));
      fail_batch(b);
    }
  }

  const clk::time_point t_end = clk::now();
  lock.lock();
  tm.submit_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t_mid - t_sub).count();
  tm.fetch_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t_end - t_mid).count();
  ++tm.batches;
  for (part* p : b->members) {
    p->state   = part::st::done;
    p->done_at = t_end;
  }
  b->state = batch::st::done;
  --inflight;
  cv.notify_all();
  return true;
}

pusch_dec_batcher::status pusch_dec_batcher::poll(part* p, unsigned wait_us)
{
  std::unique_lock<std::mutex> lock(mtx);
  if (p->state == part::st::ready) {
    try_form(lock); // may run a batch (this part's or, when more parts are ready than a batch takes, an older one)
  }
  if (p->state != part::st::done && wait_us != 0) {
    // Woken when any batch completes (this part's, or the one whose completion makes the device idle so that the next may
    // form); the time-out covers the closing deadline of the ready queue.
    cv.wait_for(lock, std::chrono::microseconds(wait_us));
    if (p->state == part::st::ready) {
      try_form(lock);
    }
  }
  if (p->state == part::st::done && !p->seen_done) {
    p->seen_done = true;
    tm.wake_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - p->done_at).count();
  }
  return (p->state == part::st::done) ? status::done : status::pending;
}

const pdc_cb_result& pusch_dec_batcher::result(const part* p, unsigned k) const
{
  static const pdc_cb_result rejected = {-1, 0, 1, 0};
  return (p->index[k] < 0) ? rejected : p->owner->results[p->index[k]];
}

const uint8_t* pusch_dec_batcher::bits(const part* p, unsigned k) const
{
  return p->owner->bits.data() + (size_t)std::max(p->index[k], 0) * BITS_STRIDE;
}

const int8_t* pusch_dec_batcher::soft(const part* p, unsigned k) const
{
  // rejected code block: nothing came back (the caller keeps its own buffer content)
  return (p->index[k] < 0) ? nullptr : p->owner->soft.data() + p->owner->descs[p->index[k]].soft_offset;
}

void pusch_dec_batcher::release(part* p)
{
  std::unique_lock<std::mutex> lock(mtx);
  if (p->state == part::st::ready) {
    // Abandoned after sealing (free_queue without dequeuing): take it out of the queue.
    ready.erase(std::remove(ready.begin(), ready.end(), p), ready.end());
  }
  while (p->state == part::st::inflight) {
    cv.wait(lock); // abandoned while its batch runs: the region must not be reused under the copy / the kernels
  }
  if (p->state == part::st::done) {
    batch* b = p->owner;
    if (++b->released == b->members.size()) {
      b->state = batch::st::idle;
      b->soft.clear();
    }
  }
  p->state = part::st::idle;
  p->owner = nullptr;
  cv_room.notify_all();
  cv.notify_all(); // a batch buffer may have become free: a ready queue that was waiting for one can form now
}

pusch_dec_batcher::timing pusch_dec_batcher::timing_so_far() const
{
  std::lock_guard<std::mutex> lock(mtx);
  timing t  = tm;
  t.add_ns  = add_ns.load();
  return t;
}

uint64_t pusch_dec_batcher::nof_batches_submitted() const
{
  std::lock_guard<std::mutex> lock(mtx);
  return stat_batches;
}

uint64_t pusch_dec_batcher::nof_code_blocks_submitted() const
{
  std::lock_guard<std::mutex> lock(mtx);
  return stat_cbs;
}

std::shared_ptr<pusch_dec_batcher> srsran::cuda::shared_batcher_of(const std::shared_ptr<pusch_dec_context>& context,
                                                                   const pusch_dec_batcher_config&           cfg)
{
  static std::mutex                                                  reg_mtx;
  static std::map<const pdc_ctx*, std::weak_ptr<pusch_dec_batcher>> registry;
  if (!context) {
    throw std::runtime_error("shared_batcher_of: an accelerator context is required");
  }
  std::lock_guard<std::mutex> lock(reg_mtx);
  for (auto it = registry.begin(); it != registry.end();) { // contexts that are gone (their address may be reused)
    it = it->second.expired() ? registry.erase(it) : std::next(it);
  }
  auto& slot = registry[context->get()];
  auto  b    = slot.lock();
  if (!b) {
    b    = std::make_shared<pusch_dec_batcher>(context, cfg);
    slot = b;
  }
  return b;
}
