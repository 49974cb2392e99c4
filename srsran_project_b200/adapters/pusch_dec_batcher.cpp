// See pusch_dec_batcher.h.
#include "pusch_dec_batcher.h"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <thread>

using namespace srsran::cuda;
using clk = std::chrono::steady_clock;

struct pusch_dec_batcher::batch {
  enum class st { idle, open, closing, inflight, fetching, done };
  st                         state = st::idle;
  std::vector<part>          parts;
  unsigned                   joined = 0, sealed = 0, released = 0;
  unsigned                   cb_reserved = 0;
  size_t                     llr_reserved = 0;
  bool                       full   = false; // a joiner did not fit: close as soon as every part is sealed
  bool                       failed = false;
  int8_t*                    llrs   = nullptr; // page-locked, batch_llrs bytes
  std::vector<pdc_cb_desc>   descs;
  std::vector<pdc_cb_result> results;
  std::vector<uint8_t>       bits;
  std::vector<int8_t>        soft;
  pdc_ticket                 ticket = 0;
  clk::time_point            opened;
};

pusch_dec_batcher::pusch_dec_batcher(std::shared_ptr<pusch_dec_context> context, const pusch_dec_batcher_config& cfg_) :
  ctx(std::move(context)), cfg(cfg_)
{
  if (!ctx) {
    throw std::runtime_error("pusch_dec_batcher: an accelerator context is required");
  }
  const pusch_dec_context_config& c = ctx->config();
  batch_cbs                         = c.max_batch_cbs;
  batch_llrs                        = c.max_batch_llrs;
  cfg.max_parts                     = std::max(1u, cfg.max_parts);
  cfg.nof_batches                   = std::max(1u, std::min(cfg.nof_batches, c.nof_instances));
  // LLRs already lie in page-locked memory when the batch is submitted: the DMA engine reads them from there.
  pdc_set_zero_copy(ctx->get(), 1);
  for (unsigned i = 0; i != cfg.nof_batches; ++i) {
    auto b = std::make_unique<batch>();
    b->parts.resize(cfg.max_parts);
    b->llrs = static_cast<int8_t*>(pdc_host_alloc(batch_llrs));
    if (!b->llrs) {
      throw std::runtime_error("pusch_dec_batcher: page-locked staging could not be allocated");
    }
    b->descs.reserve(batch_cbs);
    b->results.resize(batch_cbs);
    b->bits.resize((size_t)batch_cbs * BITS_STRIDE);
    batches.push_back(std::move(b));
  }
}

pusch_dec_batcher::~pusch_dec_batcher()
{
  for (auto& b : batches) {
    if (b->state == batch::st::inflight) {
      // Drain: the staging must not go away under a running copy.
      pdc_fetch(ctx->get(), b->ticket, b->bits.data(), b->bits.size(), b->results.data(), nullptr, 0);
    }
    pdc_host_free(b->llrs);
  }
}

pusch_dec_batcher::part* pusch_dec_batcher::join(unsigned max_cbs, size_t max_llr_bytes, unsigned wait_us)
{
  max_cbs       = std::max(1u, std::min(max_cbs, batch_cbs));
  max_llr_bytes = std::min((max_llr_bytes + 15) & ~(size_t)15, batch_llrs & ~(size_t)15);
  std::unique_lock<std::mutex> lock(mtx);
  for (int attempt = 0; attempt != 2; ++attempt) {
    part* p = join_locked(max_cbs, max_llr_bytes);
    if (p || wait_us == 0 || attempt == 1) {
      return p;
    }
    cv_room.wait_for(lock, std::chrono::microseconds(wait_us));
  }
  return nullptr;
}

pusch_dec_batcher::part* pusch_dec_batcher::join_locked(unsigned max_cbs, size_t max_llr_bytes)
{
  if (open && (open->joined == cfg.max_parts || open->cb_reserved + max_cbs > batch_cbs ||
               open->llr_reserved + max_llr_bytes > batch_llrs)) {
    open->full = true;
    open       = nullptr;
  }
  if (!open) {
    for (auto& b : batches) {
      if (b->state == batch::st::done && b->released == b->joined) {
        b->state = batch::st::idle; // wound up by its driver after every part had already been released
        b->soft.clear();
      }
      if (b->state == batch::st::idle) {
        open           = b.get();
        open->state    = batch::st::open;
        open->joined   = open->sealed = open->released = 0;
        open->cb_reserved  = 0;
        open->llr_reserved = 0;
        open->full = open->failed = false;
        open->opened = clk::now();
        break;
      }
    }
    if (!open) {
      return nullptr;
    }
  }
  part& p    = open->parts[open->joined++];
  p.owner    = open;
  p.llrs     = open->llrs + open->llr_reserved;
  p.llr_base = open->llr_reserved;
  p.llr_cap  = max_llr_bytes;
  p.llr_used = 0;
  p.cb_cap   = max_cbs;
  p.index.clear();
  p.sealed   = false;
  p.in_use   = true;
  p.descs.clear();
  p.soft.clear();
  open->cb_reserved += max_cbs;
  open->llr_reserved += max_llr_bytes;
  return &p;
}

bool pusch_dec_batcher::add(part* p, pdc_cb_desc d, const int8_t* llrs, size_t nof_llrs, const int8_t* soft_in, size_t nof_soft)
{
  // Only the owning thread touches an unsealed part: no lock.
  const size_t off = (p->llr_used + 15) & ~(size_t)15;
  if (p->sealed || p->descs.size() >= p->cb_cap || off + nof_llrs > p->llr_cap) {
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
  std::memcpy(p->llrs + off, llrs, nof_llrs);
  p->llr_used  = off + nof_llrs;
  d.llr_offset = p->llr_base + off;
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

void pusch_dec_batcher::seal(part* p)
{
  std::lock_guard<std::mutex> lock(mtx);
  if (!p->sealed) {
    p->sealed = true;
    p->owner->sealed++;
  }
}

void pusch_dec_batcher::try_advance(batch* b, std::unique_lock<std::mutex>& lock)
{
  if (b->state == batch::st::open && b->sealed == b->joined && b->joined != 0) {
    const bool timeout = clk::now() - b->opened >= std::chrono::microseconds(cfg.close_after_us);
    if (!(b->full || b->joined == cfg.max_parts || inflight == 0 || timeout)) {
      return;
    }
    b->state = batch::st::closing;
    if (open == b) {
      open = nullptr;
    }
    ++inflight;
    // One descriptor array over all parts; outputs and host soft buffers at fixed strides.
    b->descs.clear();
    size_t nsoft = 0, llr_end = 0;
    for (unsigned i = 0; i != b->joined; ++i) {
      part&  p  = b->parts[i];
      size_t ps = 0;
      for (size_t k = 0; k != p.descs.size(); ++k) {
        if (p.index[k] < 0) {
          continue;
        }
        pdc_cb_desc d = p.descs[k];
        p.index[k]    = (int)b->descs.size();
        d.out_offset  = b->descs.size() * BITS_STRIDE;
        if (d.harq_slot < 0 && !(d.flags & PDC_F_DECODE_ONLY)) {
          d.soft_offset = nsoft * SOFT_STRIDE;
          if (b->soft.size() < (nsoft + 1) * SOFT_STRIDE) {
            b->soft.resize((nsoft + 1) * SOFT_STRIDE);
          }
          std::memcpy(b->soft.data() + d.soft_offset, p.soft.data() + ps, SOFT_STRIDE);
          ps += SOFT_STRIDE;
          ++nsoft;
        }
        b->descs.push_back(d);
      }
      llr_end = std::max(llr_end, p.llr_base + p.llr_used);
    }
    const unsigned n = (unsigned)b->descs.size();
    stat_batches++;
    stat_cbs += n;
    lock.unlock();
    pdc_status st = PDC_OK;
    if (n != 0) {
      while ((st = pdc_submit(ctx->get(), b->descs.data(), n, b->llrs, llr_end, nsoft ? b->soft.data() : nullptr,
                              nsoft * SOFT_STRIDE, &b->ticket)) == PDC_ERR_CAPACITY) {
        std::this_thread::yield(); // every ticket of the context is in flight (other users of the context)
      }
    }
    lock.lock();
    if (st != PDC_OK || n == 0) {
      if (st != PDC_OK) {
        std::fprintf(stderr, "pusch_dec_batcher: batch of %u code blocks failed (status %d: %s)\n", n, (int)st,
                     pdc_last_error(ctx->get()));
      }
      for (unsigned k = 0; k != n; ++k) {
        b->results[k].iterations = -1, b->results[k].crc_pass = 0, b->results[k].status = 1, b->results[k].nof_layers = 0;
      }
      b->failed = st != PDC_OK;
      b->state  = batch::st::done;
      --inflight;
      cv.notify_all();
      return;
    }
    b->state = batch::st::inflight;
  }
  if (b->state == batch::st::inflight) {
    // The thread that gets here first waits for the device on behalf of every part of the batch (pdc_fetch blocks on the
    // ticket's event); the others sleep in poll().
    b->state = batch::st::fetching;
    lock.unlock();
    const unsigned n   = (unsigned)b->descs.size();
    const bool     any = !b->soft.empty();
    pdc_status     st  = pdc_fetch(ctx->get(), b->ticket, b->bits.data(), b->bits.size(), b->results.data(),
                              any ? b->soft.data() : nullptr, b->soft.size());
    lock.lock();
    if (st != PDC_OK) {
      std::fprintf(stderr, "pusch_dec_batcher: fetch of %u code blocks failed (status %d: %s)\n", n, (int)st,
                   pdc_last_error(ctx->get()));
      for (unsigned k = 0; k != n; ++k) {
        b->results[k].iterations = -1, b->results[k].crc_pass = 0, b->results[k].status = 1, b->results[k].nof_layers = 0;
      }
      b->failed = true;
    }
    b->state = batch::st::done;
    --inflight;
    cv.notify_all();
  }
}

pusch_dec_batcher::status pusch_dec_batcher::poll(part* p, unsigned wait_us)
{
  std::unique_lock<std::mutex> lock(mtx);
  batch*                       b = p->owner;
  if (b->state != batch::st::done) {
    try_advance(b, lock);
  }
  if (b->state != batch::st::done && wait_us != 0) {
    // Woken when any batch completes (this one, or the one whose completion lets this one close: "device idle"); the
    // time-out covers the closing deadline of an open batch.
    cv.wait_for(lock, std::chrono::microseconds(wait_us));
    if (b->state != batch::st::done) {
      try_advance(b, lock);
    }
  }
  return (b->state == batch::st::done) ? status::done : status::pending;
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
  // rejected code block: the caller's own buffer content is what it gets back (part::soft holds what it sent)
  return (p->index[k] < 0) ? nullptr : p->owner->soft.data() + p->owner->descs[p->index[k]].soft_offset;
}

void pusch_dec_batcher::release(part* p)
{
  std::lock_guard<std::mutex> lock(mtx);
  batch*                      b = p->owner;
  p->in_use                     = false;
  if (!p->sealed) { // abandoned before it was sealed (reserve_queue / free_queue without operations): counts as empty
    p->sealed = true;
    p->descs.clear();
    b->sealed++;
  }
  if (++b->released != b->joined) {
    return;
  }
  // Every part has been handed back. Normally the batch is done; parts abandoned before their results were fetched
  // (free_queue without dequeuing) leave it open or in flight: nobody will poll it again, so it is wound up here.
  if (b->state == batch::st::open) {
    if (open == b) {
      open = nullptr;
    }
  } else if (b->state == batch::st::inflight) {
    pdc_fetch(ctx->get(), b->ticket, b->bits.data(), b->bits.size(), b->results.data(), nullptr, 0);
    --inflight;
  } else if (b->state != batch::st::done) {
    return; // closing / fetching: the thread that is driving the batch finishes it; the next join finds it done + released
  }
  b->state = batch::st::idle;
  b->soft.clear();
  cv_room.notify_all();
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
