// TEST INFRASTRUCTURE. Host-only probe of the batch planner (srsran_project_b200/csrc/pdc_runtime_internal.cuh: plan_batch):
// runs the scenarios of tests/test_planner_host.py without a GPU and prints one line per planned code block:
//   scenario index list flags max_layers nz_after released_after nof_zero_slots status
// Built by the test with nvcc against libpuschdec_cuda.so (for pdc_h2_supported / pdc_h2*_smem_bytes); no CUDA call is made.
#include "pdc_runtime_internal.cuh"
#include <cstdio>
#include <vector>

static pdc_cb_desc config2_cb(uint32_t E, uint32_t Nref, int slot, bool new_data, int rv)
{
  pdc_cb_desc d;
  memset(&d, 0, sizeof(d));
  d.bg = 1, d.Z = 384, d.Qm = 8, d.rv = (uint8_t)rv, d.crc_type = PDC_CRC24B, d.max_iter = 6;
  d.flags = (uint8_t)((new_data ? PDC_F_NEW_DATA : 0) | PDC_F_EARLY_STOP);
  d.E = E, d.Nref = Nref, d.nof_filler = 16, d.harq_slot = slot;
  return d;
}

struct arena {
  std::vector<int32_t> nz;
  std::vector<uint8_t> released;
};

static void run(const char* name, std::vector<pdc_cb_desc> descs, arena& a)
{
  const uint32_t n = (uint32_t)descs.size();
  size_t         lo = 0, oo = 0;
  for (pdc_cb_desc& d : descs) {
    d.llr_offset = lo, d.out_offset = oo;
    lo += (d.E + 15) / 16 * 16, oo += 1056;
  }
  std::vector<pdc_dev_cb> dev(n);
  std::vector<uint32_t>   order(n);
  std::vector<uint8_t>    lst(n);
  std::vector<int32_t>    scratch(n), zero(n);
  uint32_t                nzero = 0;
  batch_buffers           buf;
  buf.d_llrs = reinterpret_cast<const int8_t*>(0x10000000), buf.d_bits = reinterpret_cast<uint8_t*>(0x20000000);
  buf.d_res = reinterpret_cast<pdc_cb_result*>(0x30000000), buf.d_harq = reinterpret_cast<int8_t*>(0x40000000);
  buf.d_soft = reinterpret_cast<int8_t*>(0x50000000);
  buf.harq_slots = (uint32_t)a.nz.size();
  buf.harq_nz = a.nz.data(), buf.nz_scratch = scratch.data(), buf.harq_released = a.released.data();
  buf.zero_slots = zero.data(), buf.nof_zero_slots = &nzero;
  batch_plan bp;
  pdc_status st = plan_batch(descs.data(), n, buf, dev.data(), order.data(), lst.data(), bp);
  for (uint32_t i = 0; i != n; ++i) {
    const int slot = descs[i].harq_slot;
    std::printf("%s %u %d %d %d %d %d %u %d\n", name, i, st == PDC_OK ? (int)lst[i] : -1, st == PDC_OK ? (int)dev[i].flags : -1,
                st == PDC_OK ? (int)dev[i].max_layers : -1, slot >= 0 ? a.nz[slot] : -1, slot >= 0 ? (int)a.released[slot] : -1, nzero,
                (int)st);
  }
}

int main()
{
  arena a;
  a.nz.assign(8, 0), a.released.assign(8, 0);
  // 1: limited-buffer first transmissions (Nref = 12611 < N = 25344) into empty slots; E = 8960 ends at 8976, E = 8992 at 9008
  run("clean", {config2_cb(8960, 12611, 0, true, 0), config2_cb(8992, 12611, 1, true, 0)}, a);
  // 2: new data again, lengths swapped: slot 0 now ends at 9008 (covers what it held), slot 1 at 8976 with [8976, 9008) stale
  run("swapped", {config2_cb(8992, 12611, 0, true, 0), config2_cb(8960, 12611, 1, true, 0)}, a);
  // 3: the stale stretch of slot 1 persists for a transmission of the short length, is covered again by the long one
  run("persist", {config2_cb(8960, 12611, 1, true, 0)}, a);
  run("covered", {config2_cb(8992, 12611, 1, true, 0)}, a);
  // 4: a retransmission (rv 2) may leave soft bits anywhere: the bound is the whole buffer; new data then takes the general list
  run("retx", {config2_cb(8960, 12611, 2, true, 0)}, a);
  run("retx", {config2_cb(8960, 12611, 2, false, 2)}, a);
  run("after_retx", {config2_cb(8960, 12611, 2, true, 0)}, a);
  // 5: released (pdc_harq_release): reads as zeros; a first transmission rewrites everything, nothing to clear ...
  a.released[2] = 1;
  run("released_first_tx", {config2_cb(8960, 12611, 2, true, 0)}, a);
  // ... a combining transmission into a released slot has it cleared first
  a.nz[3] = 25344, a.released[3] = 1;
  run("released_retx", {config2_cb(8960, 12611, 3, false, 2)}, a);
  // 6: without a buffer limit the whole buffer is rewritten whatever the slot held
  a.nz[4] = 25344;
  run("no_limit", {config2_cb(8960, 0, 4, true, 0)}, a);
  // 7: a rejected batch (second descriptor invalid: Z = 385) changes nothing
  a.nz[5] = 777;
  pdc_cb_desc bad = config2_cb(8960, 12611, 6, true, 0);
  bad.Z           = 385;
  run("rejected", {config2_cb(8960, 12611, 5, true, 0), bad}, a);
  // 8: host-exchanged soft buffers (harq_slot < 0) are never tracked: limited-buffer code blocks take the general list
  run("host_soft", {config2_cb(8960, 12611, -1, true, 0)}, a);
  return 0;
}
