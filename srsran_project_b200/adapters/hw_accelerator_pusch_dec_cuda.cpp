// See hw_accelerator_pusch_dec_cuda.h. One object serves one transport block at a time (the reference takes it from a
// bounded pool per TB, pusch_decoder_hw_impl.cpp:151, factories.cpp:131-138); the objects of one factory share the GPU
// context(s) and are driven concurrently from the PUSCH decoder executor threads.
//
// Protocol as driven by pusch_decoder_hw_impl::run_asynch_hw_decoder (pusch_decoder_hw_impl.cpp:210-360):
//   reserve_queue -> { configure_operation(cb) ; enqueue_operation(cb) } for every code block still to decode
//                 -> { dequeue_operation(cb) ; read_operation_outputs(cb) } for the same code blocks -> free_queue.
// enqueue copies the code block's LLRs straight into the page-locked staging of the batch that is open on the code block's
// GPU (pusch_dec_batcher.h) — the batch is shared with every other accelerator object of the factory; the first dequeue
// seals this object's share. dequeue returns false ("not ready", the reference spins on it: :325-338) until the batch has
// run: the GPU sees ONE submission for the transport blocks of all objects that were enqueued while it was busy.
// Errors never throw: a failed submission reports every code block of the batch as "CRC failed".
#include "hw_accelerator_pusch_dec_cuda.h"
#include "srsran/ran/sch/modulation_scheme.h"
#include <array>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

using namespace srsran;
using namespace srsran::hal;

namespace {

constexpr unsigned MAX_CBS  = 152; // ldpc::MAX_NOF_SEGMENTS
constexpr unsigned MAX_GPUS = 16;

using batcher = cuda::pusch_dec_batcher;

class hw_accelerator_pusch_dec_cuda : public hw_accelerator_pusch_dec
{
public:
  hw_accelerator_pusch_dec_cuda(std::vector<std::shared_ptr<batcher>> batchers_, bool external_, bool generic_, unsigned id_) :
    batchers(std::move(batchers_)), external(external_), generic(generic_), instance_id(id_)
  {
    parts.fill(nullptr);
    for (auto& s : where) {
      s = {-1, 0};
    }
  }
  ~hw_accelerator_pusch_dec_cuda() override { drop_parts(); }

  void reserve_queue() override { reset(); }
  void free_queue() override { reset(); }

  void configure_operation(const hw_pusch_decoder_configuration& config, unsigned cb_index) override
  {
    if (cb_index < MAX_CBS) {
      cfgs[cb_index] = config;
    }
  }

  bool enqueue_operation(span<const int8_t> data, span<const int8_t> aux_data, unsigned cb_index) override
  {
    if (cb_index >= MAX_CBS || sealed) {
      return false; // the previous share has not been drained yet: the caller dequeues and retries
    }
    const hw_pusch_decoder_configuration& c = cfgs[cb_index];
    const unsigned                        G = static_cast<unsigned>(batchers.size());
    // Placement: by absolute code-block identifier, so that every transmission of a code block meets its soft buffer.
    const unsigned g = external ? (c.absolute_cb_id % G) : (instance_id % G);
    if (!parts[g]) {
      // Room for this GPU's share of the transport block (code blocks are spread evenly: ids are handed out in order).
      const unsigned nof_cbs = std::max(1u, std::min(c.nof_segments, MAX_CBS));
      const unsigned share   = external ? (nof_cbs + G - 1) / G + 1 : 1;
      parts[g]               = batchers[g]->join(share, (size_t)share * ((data.size() + 63 + 32) & ~(size_t)63), 200);
      if (!parts[g]) {
        return false; // every staging buffer of that GPU is taken: back-pressure, the caller retries
      }
      counts[g] = 0;
    }
    pdc_cb_desc d = {};
    d.bg          = (c.base_graph_index == ldpc_base_graph_type::BG1) ? 1 : 2;
    d.Qm          = get_bits_per_symbol(c.modulation);
    d.rv          = c.rv;
    d.crc_type    = (c.cb_crc_type == hw_dec_cb_crc_type::CRC16)
                        ? PDC_CRC16
                        : ((c.cb_crc_type == hw_dec_cb_crc_type::CRC24A) ? PDC_CRC24A : PDC_CRC24B);
    d.max_iter    = (c.max_nof_ldpc_iterations > 255) ? 255 : c.max_nof_ldpc_iterations;
    d.flags       = (c.new_data ? PDC_F_NEW_DATA : 0) | (c.use_early_stop ? PDC_F_EARLY_STOP : 0) |
              (generic ? PDC_F_GENERIC : 0);
    d.Z          = c.lifting_size;
    d.E          = data.size();
    d.Nref       = c.Nref;
    d.nof_filler = c.nof_filler_bits;
    // External HARQ: arena slot on the code block's GPU. Otherwise the host buffer travels with the operation.
    d.harq_slot = external ? static_cast<int32_t>(c.absolute_cb_id / G) : -1;
    if (!batchers[g]->add(parts[g], d, data.data(), data.size(), external ? nullptr : aux_data.data(),
                          external ? 0 : aux_data.size())) {
      return false; // this share is full: the caller drains it, then enqueues the rest
    }
    where[cb_index] = {static_cast<int>(g), counts[g]++};
    ++nof_enqueued;
    return true;
  }

  bool dequeue_operation(span<uint8_t> data, span<int8_t> aux_data, unsigned segment_index) override
  {
    if (segment_index >= MAX_CBS || where[segment_index].gpu < 0) {
      return false;
    }
    if (!sealed) {
      for (unsigned g = 0; g != batchers.size(); ++g) {
        if (parts[g]) {
          batchers[g]->seal(parts[g]);
        }
      }
      sealed = true;
    }
    const slot     s = where[segment_index];
    batcher&       b = *batchers[s.gpu];
    batcher::part* p = parts[s.gpu];
    // One poll (a mutex) per share, not per code block: once the share's batch is done it stays done until released.
    if (!done[s.gpu]) {
      if (b.poll(p, 100) != batcher::status::done) {
        return false; // not ready (after a short sleep: the caller spins on this call)
      }
      done[s.gpu] = true;
    }
    const pdc_cb_result& r = b.result(p, s.index);
    const pdc_cb_desc&   d = p->descs[s.index];
    const unsigned       K = pdc_cb_message_bits(d.bg, d.Z);
    std::memcpy(data.data(), b.bits(p, s.index), std::min<size_t>(data.size(), (K + 7) / 8));
    if (!external && !aux_data.empty()) {
      std::memcpy(aux_data.data(), b.soft(p, s.index), std::min<size_t>(aux_data.size(), batcher::SOFT_STRIDE));
    }
    outputs[segment_index].CRC_pass            = (r.crc_pass != 0);
    outputs[segment_index].nof_ldpc_iterations = (r.iterations > 0) ? r.iterations : d.max_iter;
    where[segment_index].gpu                   = -1;
    if (++nof_dequeued == nof_enqueued) {
      drop_parts(); // everything handed back: the object accepts a new share (outputs[] stay readable)
    }
    return true;
  }

  void read_operation_outputs(hw_pusch_decoder_outputs& out, unsigned cb_index, unsigned /*absolute_cb_id*/) override
  {
    out = (cb_index < MAX_CBS) ? outputs[cb_index] : hw_pusch_decoder_outputs{false, 0};
  }

  // "Frees the HARQ context entry" (hw_accelerator_pusch_dec.h:107-110; called per code block once its transport block has
  // passed the CRC, pusch_decoder_hw_impl.cpp:413-428): the arena slot reads as zeros from now on. No CUDA call: the runtime
  // notes it and clears lazily (pdc_harq_release). An entry that is never freed (a transport block that failed for good)
  // keeps its content, and a later first transmission meets what the reference's dematcher would meet in a host buffer that
  // rx_buffer_codeblock_pool hands out again without clearing: whatever the write set of the new transmission leaves.
  void free_harq_context_entry(unsigned absolute_cb_id) override
  {
    if (external) {
      const unsigned G = static_cast<unsigned>(batchers.size());
      pdc_harq_release(batchers[absolute_cb_id % G]->context().get(), static_cast<int32_t>(absolute_cb_id / G));
    }
  }

  bool is_harq_external() const override { return external; }

private:
  struct slot {
    int      gpu;
    unsigned index;
  };

  void drop_parts()
  {
    for (unsigned g = 0; g != batchers.size(); ++g) {
      if (parts[g]) {
        batchers[g]->release(parts[g]);
        parts[g] = nullptr;
      }
      done[g] = false;
    }
    nof_enqueued = nof_dequeued = 0;
    sealed                      = false;
  }

  void reset()
  {
    drop_parts();
    for (auto& s : where) {
      s.gpu = -1;
    }
  }

  std::vector<std::shared_ptr<batcher>> batchers; // one per GPU
  bool                                  external;
  bool                                  generic;
  unsigned                              instance_id;

  std::array<hw_pusch_decoder_configuration, MAX_CBS> cfgs    = {};
  std::array<hw_pusch_decoder_outputs, MAX_CBS>       outputs = {};
  std::array<slot, MAX_CBS>                           where;
  std::array<batcher::part*, MAX_GPUS>                parts;
  std::array<unsigned, MAX_GPUS>                      counts = {};
  std::array<bool, MAX_GPUS>                          done   = {};
  unsigned                                            nof_enqueued = 0, nof_dequeued = 0;
  bool                                                sealed = false; // this object's share is closed: results are awaited
};

class hw_accelerator_pusch_dec_factory_cuda : public hw_accelerator_pusch_dec_factory
{
public:
  explicit hw_accelerator_pusch_dec_factory_cuda(const cuda_hwacc_pusch_dec_factory_configuration& cfg_) : cfg(cfg_)
  {
    std::vector<std::shared_ptr<cuda::pusch_dec_context>> contexts;
    if (cfg.context) {
      contexts.push_back(cfg.context);
    } else {
      std::vector<int> devices = cfg.devices.empty() ? std::vector<int>{0} : cfg.devices;
      if (devices.size() > MAX_GPUS) {
        throw std::runtime_error("cuda pusch_dec accelerator: too many devices");
      }
      const unsigned G = static_cast<unsigned>(devices.size());
      for (int dev : devices) {
        // Sized for the batcher: max_parts transport blocks of the largest codeword per submission.
        cuda::pusch_dec_context_config c;
        c.device          = dev;
        c.max_batch_cbs   = std::max(1u, cfg.batching.max_parts) * MAX_CBS;
        c.max_batch_llrs  = (size_t)std::max(1u, cfg.batching.max_parts) * (1536u << 10);
        c.harq_codeblocks = cfg.external_harq ? (cfg.harq_context_size + G - 1) / G : 0;
        c.nof_instances   = std::max(1u, cfg.batching.nof_batches);
        contexts.push_back(std::make_shared<cuda::pusch_dec_context>(c));
      }
    }
    for (auto& c : contexts) {
      batchers.push_back(cuda::shared_batcher_of(c, cfg.batching));
    }
  }
  std::unique_ptr<hw_accelerator_pusch_dec> create() override
  {
    return std::make_unique<hw_accelerator_pusch_dec_cuda>(batchers, cfg.external_harq, cfg.generic_arithmetic, next_id++);
  }
  void timing(uint64_t out[8]) const
  {
    std::fill(out, out + 8, 0);
    for (auto& b : batchers) {
      const batcher::timing t = b->timing_so_far();
      out[0] += t.parts, out[1] += t.batches, out[2] += t.queue_ns, out[3] += t.submit_ns, out[4] += t.fetch_ns;
      out[5] += t.wake_ns, out[6] += t.join_wait_ns, out[7] += t.add_ns;
    }
  }
  void stats(uint64_t& nb, uint64_t& ncb) const
  {
    nb = ncb = 0;
    for (auto& b : batchers) {
      nb += b->nof_batches_submitted();
      ncb += b->nof_code_blocks_submitted();
    }
  }

private:
  cuda_hwacc_pusch_dec_factory_configuration cfg;
  std::vector<std::shared_ptr<batcher>>      batchers;
  std::atomic<unsigned>                      next_id{0};
};

} // namespace

std::shared_ptr<hw_accelerator_pusch_dec_factory>
srsran::hal::create_cuda_pusch_dec_acc_factory(const cuda_hwacc_pusch_dec_factory_configuration& accelerator_config)
{
  return std::make_shared<hw_accelerator_pusch_dec_factory_cuda>(accelerator_config);
}

void srsran::hal::cuda_pusch_dec_acc_factory_timing(const hw_accelerator_pusch_dec_factory& factory, uint64_t out[8])
{
  std::fill(out, out + 8, 0);
  if (auto* f = dynamic_cast<const hw_accelerator_pusch_dec_factory_cuda*>(&factory)) {
    f->timing(out);
  }
}

void srsran::hal::cuda_pusch_dec_acc_factory_stats(const hw_accelerator_pusch_dec_factory& factory, uint64_t& batches,
                                                  uint64_t& code_blocks)
{
  batches = code_blocks = 0;
  if (auto* f = dynamic_cast<const hw_accelerator_pusch_dec_factory_cuda*>(&factory)) {
    f->stats(batches, code_blocks);
  }
}
