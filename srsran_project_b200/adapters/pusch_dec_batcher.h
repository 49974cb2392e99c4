// Cross-instance batching of code-block operations in front of one accelerator context (one GPU).
//
// The reference drives its accelerator objects per transport block and serially per object: nof_pusch_decoder_threads
// objects "shared for all the PUSCH decoders" (lib/phy/upper/channel_processors/pusch/factories.cpp:130-138), each running
// pusch_decoder_hw_impl::run_asynch_hw_decoder (pusch_decoder_hw_impl.cpp:149-364) on its own executor thread; the software
// operators are pooled the same way (factories.cpp:82-94). One GPU launch per object and transport block leaves the device
// mostly idle (152 thread blocks on a machine that holds 592) and pays one PCIe round trip per transport block. The batcher
// coalesces the operations of ALL objects that share a context:
//
//   join()     an object takes a PART: a region of the context's staging arena (page-locked host memory mirrored by device
//              memory, pdc_stage_*) for one transport block, or one code block of the ldpc_decoder / ldpc_rate_dematcher operators
//   add()      the object's code blocks: LLRs are copied straight into the page-locked region (the only host copy before
//              the DMA), descriptors are collected per part
//   seal()     no more code blocks will follow: the region is pushed to the device at once (asynchronous copy on the arena's
//              stream) and the part joins the queue of ready parts — it does not wait for anybody else's part
//   poll()     "not ready" until the part's batch has run (hw_accelerator::dequeue_operation may return false:
//              pusch_decoder_hw_impl.cpp:325-338 spins on it). A batch is formed from the ready parts — ONE pdc_submit_staged:
//              one launch per kernel variant over LLRs that are already on the device, results written straight to page-locked memory — as soon as max_parts are
//              ready, OR fewer than eager_inflight batches are running, OR the oldest ready part has waited close_after_us: a device with room is
//              never kept waiting, a busy one collects work meanwhile. The thread whose poll forms the batch submits it and
//              waits for the device on behalf of everybody; the others sleep on a condition variable.
//   release()  results handed back; the region and (once all its parts are back) the batch's result buffers are reused
//
// Thread-safe. No CPU fallback: a failed submission marks every code block of the batch "CRC failed" (status 1).
#pragma once

#include "cuda_pusch_dec_context.h"
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <vector>

namespace srsran {
namespace cuda {

struct pusch_dec_batcher_config {
  /// Parts (transport blocks) per batch. Clamped to what the context's batch capacity allows.
  unsigned max_parts = 8;
  /// A batch that is neither full nor facing an idle device closes after this long.
  unsigned close_after_us = 150;
  /// Batches whose results may be outstanding at once (result buffers); at most the context's nof_instances.
  unsigned nof_batches = 8;
  /// A part's region travels to the device in pieces of this size while its owner is still adding code blocks, so that only
  /// the last piece is on the road when the part is sealed. 0: one transfer at seal().
  size_t push_chunk_bytes = 0; // measured with 448 KB pieces on config 2: no gain (more entries on the copy stream)
  /// A new batch forms at once while fewer than this many are running on the device ("the device is idle enough"); beyond
  /// that, parts wait for max_parts or close_after_us.
  unsigned eager_inflight = 6;
};

class pusch_dec_batcher
{
public:
  static constexpr unsigned MAX_PART_CBS = 152;      // ldpc::MAX_NOF_SEGMENTS
  static constexpr size_t   SOFT_STRIDE  = 66 * 384; // largest soft buffer
  static constexpr size_t   BITS_STRIDE  = 8448 / 8; // largest message, bytes

  struct batch;
  struct part {
    enum class st { idle, filling, ready, inflight, done };
    st                       state = st::idle;
    batch*                   owner = nullptr; // the batch that carries the part (inflight, done)
    int8_t*                  llrs  = nullptr; // this part's region of the staging arena (host side)
    size_t                   llr_base = 0, llr_cap = 0, llr_used = 0, llr_pushed = 0;
    unsigned                 cb_cap = 0;
    std::vector<pdc_cb_desc> descs;  // llr offsets relative to the arena; out / soft offsets are assigned when the batch forms
    std::vector<int8_t>      soft;   // host-side HARQ buffers travelling with the operations (harq_slot < 0), SOFT_STRIDE each
    std::vector<int>         index;  // per code block: position in the batch once it has formed, -1 = rejected by add()
    std::chrono::steady_clock::time_point sealed_at, done_at;
    bool                     seen_done = false;
  };
  /// Where the time of the parts and batches went so far, nanoseconds summed over parts / batches (for the bench and tuning).
  struct timing {
    uint64_t parts = 0, batches = 0;
    uint64_t queue_ns  = 0; // sealed -> taken into a batch, per part
    uint64_t submit_ns = 0; // pdc_submit_staged, per batch
    uint64_t fetch_ns  = 0; // pdc_fetch (wait for the device + copy out), per batch
    uint64_t wake_ns   = 0; // batch done -> the part's owner sees it, per part
    uint64_t join_wait_ns = 0; // join() sleeping for a free region
    uint64_t add_ns    = 0; // add(): the copy into the page-locked region
  };
  enum class status { pending, done };

  pusch_dec_batcher(std::shared_ptr<pusch_dec_context> context, const pusch_dec_batcher_config& cfg);
  ~pusch_dec_batcher();
  pusch_dec_batcher(const pusch_dec_batcher&)            = delete;
  pusch_dec_batcher& operator=(const pusch_dec_batcher&) = delete;

  /// A part of the open batch with room for max_cbs code blocks and max_llr_bytes of LLRs (each code block 64-byte aligned),
  /// or nullptr when every batch buffer is taken (try again: back-pressure, not an error). A request larger than a whole
  /// batch is cut down to the batch: add() then reports "full" and the caller continues with another part.
  /// wait_us > 0: when no batch buffer is free the call sleeps up to that long for one (woken when a batch is recycled).
  part* join(unsigned max_cbs, size_t max_llr_bytes, unsigned wait_us = 0);
  /// Appends one code block. d.llr_offset / out_offset / soft_offset are assigned here. soft_in: the code block's host
  /// soft buffer (harq_slot < 0 and not decode-only), may be empty. Returns false if the part is full.
  bool add(part* p, pdc_cb_desc d, const int8_t* llrs, size_t nof_llrs, const int8_t* soft_in, size_t nof_soft);
  void seal(part* p);
  /// wait_us > 0: when the batch is not done yet the call may sleep up to that long (woken as soon as it is) instead of
  /// returning at once: callers that spin on the result (the reference's dequeue loop) then leave the cores to the threads
  /// that still have host work to do.
  status poll(part* p, unsigned wait_us = 0);
  /// After status::done: outcome of the part's k-th code block.
  const pdc_cb_result& result(const part* p, unsigned k) const;
  const uint8_t*       bits(const part* p, unsigned k) const;
  const int8_t*        soft(const part* p, unsigned k) const; // updated host soft buffer (harq_slot < 0)
  void                 release(part* p);

  pusch_dec_context& context() { return *ctx; }
  /// Batches submitted and code blocks carried so far (for tests and the bench: mean batch size = code_blocks / batches).
  uint64_t nof_batches_submitted() const;
  uint64_t nof_code_blocks_submitted() const;
  timing   timing_so_far() const;

private:
  bool  try_form(std::unique_lock<std::mutex>& lock);
  part* join_locked(unsigned max_cbs, size_t max_llr_bytes);
  void  fail_batch(batch* b);
  void  push_pending(part* p);

  std::shared_ptr<pusch_dec_context>  ctx;
  pusch_dec_batcher_config            cfg;
  mutable std::mutex                  mtx;
  std::condition_variable             cv;      // a batch completed (or failed): waiters re-evaluate
  std::condition_variable             cv_room; // a part region or a batch was recycled: joiners re-evaluate
  pdc_stage*                          stage = nullptr;
  std::vector<part>                   parts;   // one per region of the arena
  std::vector<part*>                  ready;   // sealed parts waiting for a batch, oldest first
  std::vector<std::unique_ptr<batch>> batches;
  unsigned                            inflight = 0;
  uint64_t                            stat_batches = 0, stat_cbs = 0;
  timing                              tm;
  std::atomic<uint64_t>               add_ns{0};
  unsigned                            batch_cbs = 0;
  size_t                              part_bytes = 0;
};

/// The batcher of a context: every adapter object that shares the context shares its batcher (created on first use with
/// `cfg`; later calls return the same object whatever their cfg).
std::shared_ptr<pusch_dec_batcher> shared_batcher_of(const std::shared_ptr<pusch_dec_context>& context,
                                                     const pusch_dec_batcher_config&           cfg = {});

} // namespace cuda
} // namespace srsran
