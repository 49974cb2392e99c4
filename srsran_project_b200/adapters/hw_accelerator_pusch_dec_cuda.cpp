// See hw_accelerator_pusch_dec_cuda.h. One object serves one transport block at a time (the reference takes it from a
// bounded pool per TB, pusch_decoder_hw_impl.cpp:151, factories.cpp:131-138); objects of one factory share the GPU
// context and are driven concurrently from the PUSCH decoder executor threads.
//
// Protocol as driven by pusch_decoder_hw_impl::run_asynch_hw_decoder (pusch_decoder_hw_impl.cpp:210-360):
//   reserve_queue -> { configure_operation(cb) ; enqueue_operation(cb) } for every code block still to decode
//                 -> { dequeue_operation(cb) ; read_operation_outputs(cb) } for the same code blocks -> free_queue.
// enqueue only records the code block (descriptor + LLRs into one contiguous staging vector); the first dequeue submits
// everything recorded so far as ONE batch (one H2D copy, one kernel launch of one CTA per code block, one D2H copy) and
// waits for it. With external HARQ the caller enqueues the whole TB before it dequeues, so a TB is one launch.
// Errors never throw: a failed submission reports every code block of the batch as "CRC failed".
#include "hw_accelerator_pusch_dec_cuda.h"
#include "srsran/ran/sch/modulation_scheme.h"
#include <array>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

using namespace srsran;
using namespace srsran::hal;

namespace {

constexpr unsigned MAX_CBS     = 152;        // ldpc::MAX_NOF_SEGMENTS
constexpr unsigned SOFT_STRIDE = 66 * 384;   // largest soft buffer, one HARQ arena slot
constexpr unsigned BITS_STRIDE = 8448 / 8;   // largest message, bytes

class hw_accelerator_pusch_dec_cuda : public hw_accelerator_pusch_dec
{
public:
  hw_accelerator_pusch_dec_cuda(std::shared_ptr<cuda::pusch_dec_context> context_, bool external_, bool generic_) :
    context(std::move(context_)), external(external_), generic(generic_)
  {
    max_cbs = std::min<unsigned>(MAX_CBS, context->config().max_batch_cbs);
    descs.reserve(max_cbs);
    llrs.reserve(context->config().max_batch_llrs);
    bits.resize((size_t)max_cbs * BITS_STRIDE);
    results.resize(max_cbs);
    slot_of_cb.fill(-1);
  }

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
    if (cb_index >= MAX_CBS || ready || descs.size() >= max_cbs) {
      return false; // batch full, or the previous batch has not been drained yet: the caller dequeues and retries
    }
    const hw_pusch_decoder_configuration& c = cfgs[cb_index];
    const size_t llr_off = (llrs.size() + 15) / 16 * 16;
    if (llr_off + data.size() > context->config().max_batch_llrs) {
      return false;
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
    d.Z           = c.lifting_size;
    d.E           = data.size();
    d.Nref        = c.Nref;
    d.nof_filler  = c.nof_filler_bits;
    d.llr_offset  = llr_off;
    d.out_offset  = descs.size() * BITS_STRIDE;
    d.soft_offset = 0;
    if (external) {
      d.harq_slot = static_cast<int32_t>(c.absolute_cb_id);
    } else {
      // Host-side HARQ buffer: it travels with the operation and comes back updated in dequeue_operation.
      d.harq_slot   = -1;
      d.soft_offset = descs.size() * (size_t)SOFT_STRIDE;
      soft.resize(d.soft_offset + SOFT_STRIDE, 0);
      std::memcpy(soft.data() + d.soft_offset, aux_data.data(), std::min<size_t>(aux_data.size(), SOFT_STRIDE));
    }
    llrs.resize(llr_off + data.size());
    std::memcpy(llrs.data() + llr_off, data.data(), data.size());
    slot_of_cb[cb_index] = static_cast<int>(descs.size());
    descs.push_back(d);
    return true;
  }

  bool dequeue_operation(span<uint8_t> data, span<int8_t> aux_data, unsigned segment_index) override
  {
    if (segment_index >= MAX_CBS || slot_of_cb[segment_index] < 0) {
      return false;
    }
    if (!ready) {
      run_batch();
    }
    const unsigned       k = slot_of_cb[segment_index];
    const pdc_cb_result& r = results[k];
    const unsigned       K = pdc_cb_message_bits(descs[k].bg, descs[k].Z);
    std::memcpy(data.data(), bits.data() + (size_t)k * BITS_STRIDE, std::min<size_t>(data.size(), (K + 7) / 8));
    if (!external && !aux_data.empty()) {
      std::memcpy(aux_data.data(), soft.data() + (size_t)k * SOFT_STRIDE, std::min<size_t>(aux_data.size(), SOFT_STRIDE));
    }
    outputs[segment_index].CRC_pass            = (r.crc_pass != 0);
    outputs[segment_index].nof_ldpc_iterations = (r.iterations > 0) ? r.iterations : descs[k].max_iter;
    slot_of_cb[segment_index]                  = -1;
    if (++nof_dequeued == descs.size()) {
      clear_batch(); // everything handed back: the object accepts a new batch (outputs[] stay readable)
    }
    return true;
  }

  void read_operation_outputs(hw_pusch_decoder_outputs& out, unsigned cb_index, unsigned /*absolute_cb_id*/) override
  {
    out = (cb_index < MAX_CBS) ? outputs[cb_index] : hw_pusch_decoder_outputs{false, 0};
  }

  // The arena slot of a released code block needs no clean-up: a new transmission (new_data) rewrites it with the
  // dematcher's write set, exactly like the host buffer that rx_buffer_codeblock_pool hands out again without clearing.
  void free_harq_context_entry(unsigned /*absolute_cb_id*/) override {}

  bool is_harq_external() const override { return external; }

private:
  void run_batch()
  {
    pdc_ctx*   ctx    = context->get();
    pdc_ticket ticket = 0;
    pdc_status st;
    const int8_t* soft_in = external ? nullptr : soft.data();
    while ((st = pdc_submit(ctx, descs.data(), descs.size(), llrs.data(), llrs.size(), soft_in, soft.size(), &ticket)) ==
           PDC_ERR_CAPACITY) {
      std::this_thread::yield(); // every submission slot is in use by other accelerator instances
    }
    if (st == PDC_OK) {
      st = pdc_fetch(ctx, ticket, bits.data(), bits.size(), results.data(), external ? nullptr : soft.data(), soft.size());
    }
    if (st != PDC_OK) {
      std::fprintf(stderr, "hw_accelerator_pusch_dec_cuda: batch of %zu code blocks failed (status %d: %s)\n",
                   descs.size(), (int)st, pdc_last_error(ctx));
      for (size_t k = 0; k != descs.size(); ++k) {
        results[k].iterations = -1, results[k].crc_pass = 0, results[k].status = 1, results[k].nof_layers = 0;
      }
      std::memset(bits.data(), 0, bits.size());
    }
    ready = true;
  }

  void clear_batch()
  {
    descs.clear();
    llrs.clear();
    soft.clear();
    nof_dequeued = 0;
    ready        = false;
  }

  void reset()
  {
    clear_batch();
    slot_of_cb.fill(-1);
  }

  std::shared_ptr<cuda::pusch_dec_context> context;
  bool                                     external;
  bool                                     generic;
  unsigned                                 max_cbs = MAX_CBS;

  std::array<hw_pusch_decoder_configuration, MAX_CBS> cfgs    = {};
  std::array<hw_pusch_decoder_outputs, MAX_CBS>       outputs = {};
  std::array<int, MAX_CBS>                            slot_of_cb;
  std::vector<pdc_cb_desc>                            descs;
  std::vector<int8_t>                                 llrs;
  std::vector<int8_t>                                 soft;
  std::vector<uint8_t>                                bits;
  std::vector<pdc_cb_result>                          results;
  size_t                                              nof_dequeued = 0;
  bool                                                ready        = false; // results of the current batch are available
};

class hw_accelerator_pusch_dec_factory_cuda : public hw_accelerator_pusch_dec_factory
{
public:
  explicit hw_accelerator_pusch_dec_factory_cuda(const cuda_hwacc_pusch_dec_factory_configuration& cfg_) : cfg(cfg_)
  {
    if (!cfg.context) {
      cuda::pusch_dec_context_config c;
      c.harq_codeblocks = cfg.external_harq ? 16 * MAX_CBS : 0;
      cfg.context       = std::make_shared<cuda::pusch_dec_context>(c);
    }
  }
  std::unique_ptr<hw_accelerator_pusch_dec> create() override
  {
    return std::make_unique<hw_accelerator_pusch_dec_cuda>(cfg.context, cfg.external_harq, cfg.generic_arithmetic);
  }

private:
  cuda_hwacc_pusch_dec_factory_configuration cfg;
};

} // namespace

std::shared_ptr<hw_accelerator_pusch_dec_factory>
srsran::hal::create_cuda_pusch_dec_acc_factory(const cuda_hwacc_pusch_dec_factory_configuration& accelerator_config)
{
  return std::make_shared<hw_accelerator_pusch_dec_factory_cuda>(accelerator_config);
}
