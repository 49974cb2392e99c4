// Host-side C++ adapters that put the CUDA code-block chain (libpuschdec_cuda.so, include/pusch_dec_cuda.h) behind
// srsRAN's own operator interfaces. This header: the shared accelerator context.
//
// Everything in srsran_project_b200/adapters/ is compiled against the reference's PUBLIC headers only
// (-I<srsRAN>/include); nothing from the reference's lib/ is needed. See INTEGRATION.md for where a srsRAN maintainer
// plugs these in (channel_coding_factories.cpp:114-137, upper_phy_factories.cpp:758-798, du_low_hal_factory.cpp:84-101).
#pragma once

#include "../../include/pusch_dec_cuda.h"
#include <memory>
#include <stdexcept>
#include <string>

namespace srsran {
namespace cuda {

/// Configuration of one accelerator context (one GPU).
struct pusch_dec_context_config {
  /// CUDA device ordinal.
  int device = 0;
  /// Largest number of code blocks in one submission (a transport block has at most 152, ldpc.h MAX_NOF_SEGMENTS).
  unsigned max_batch_cbs = 152;
  /// Largest number of LLR bytes in one submission (default: one 273-PRB / 4-layer / 256QAM codeword with margin).
  size_t max_batch_llrs = 2u << 20;
  /// Capacity of the device-resident HARQ soft-buffer arena, in code blocks. It is indexed by the absolute code-block
  /// identifier of the rx_buffer pool (rx_buffer_impl.h:170-177), so it must cover rx_buffer_pool_config::nof_codeblocks.
  unsigned harq_codeblocks = 0;
  /// Number of submissions that may be in flight at once: one per concurrently used decoder / accelerator instance
  /// (nof_pusch_decoder_threads in the reference's pools, factories.cpp:82-94,131-138).
  unsigned nof_instances = 4;
};

/// RAII owner of a pdc_ctx. Shared by every decoder / dematcher / accelerator instance created from one factory.
class pusch_dec_context
{
public:
  explicit pusch_dec_context(const pusch_dec_context_config& cfg_) : cfg(cfg_)
  {
    pdc_config c;
    c.device         = cfg.device;
    c.max_batch_cbs  = cfg.max_batch_cbs;
    c.max_batch_llrs = cfg.max_batch_llrs;
    c.harq_slots     = cfg.harq_codeblocks;
    c.nof_tickets    = cfg.nof_instances;
    pdc_status st    = pdc_create(&c, &ctx);
    if (st != PDC_OK) {
      // No CPU fallback: a factory asked for "cuda" on a host without a usable device must fail loudly.
      throw std::runtime_error(std::string("pdc_create failed: ") + pdc_last_error(nullptr));
    }
  }
  ~pusch_dec_context() { pdc_destroy(ctx); }
  pusch_dec_context(const pusch_dec_context&)            = delete;
  pusch_dec_context& operator=(const pusch_dec_context&) = delete;

  pdc_ctx*                        get() const { return ctx; }
  const pusch_dec_context_config& config() const { return cfg; }

private:
  pusch_dec_context_config cfg;
  pdc_ctx*                 ctx = nullptr;
};

} // namespace cuda
} // namespace srsran
