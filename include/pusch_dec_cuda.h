/*
 * pusch_dec_cuda.h — C ABI of the B200-native PUSCH code-block decode chain (libpuschdec_cuda.so).
 *
 * One code block (CB) = LDPC rate dematching with HARQ soft combining -> layered normalised min-sum LDPC decoding
 * (BG1/BG2, every lifting size) -> CB CRC with early stop, executed by one CUDA thread block on an sm_100a device.
 * The entry points are what the reference's C++ operator interfaces for this path bind to (see INTEGRATION.md for
 * the adapter classes):
 *
 *   srsran::ldpc_decoder::decode                 include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:37-73
 *   srsran::ldpc_rate_dematcher::rate_dematch    include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:35-57
 *   srsran::crc_calculator (CB CRC, early stop)  include/srsran/phy/upper/channel_coding/crc_calculator.h:31-87
 *   srsran::pusch_codeblock_decoder::decode      lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp:35-74
 *   srsran::hal::hw_accelerator_pusch_dec        include/srsran/hal/phy/upper/channel_processors/pusch/
 *                                                hw_accelerator_pusch_dec.h:83-115 (enqueue/dequeue/read_outputs/HARQ)
 *   srsran::pusch_decoder_impl::join_and_notify  lib/phy/upper/channel_processors/pusch/pusch_decoder_impl.cpp:396-505
 *     + concatenate_codeblocks                   (transport-block assembly and TB CRC24A: the pdc_*_tbs entry points)
 *   srsran::demodulation_mapper::demodulate_soft include/srsran/phy/upper/channel_modulation/demodulation_mapper.h:66-69
 *     + revert_scrambling                        lib/phy/upper/channel_processors/pusch/pusch_demodulator_impl.cpp:60-222,
 *                                                :261-262,:397-405 (the LLR producers: the pdc_demodulate* entry points)
 *   srsran::pdsch_encoder::encode                include/srsran/phy/upper/channel_processors/pdsch/pdsch_encoder.h:73
 *                                                (the downlink mirror: the pdc_pdsch_encode* entry points)
 *
 * Conventions: plain pointers and sizes only; every function returns a pdc_status (0 = ok) and never throws or
 * aborts; a context is thread-safe; there is NO CPU fallback — without a CUDA device pdc_create fails.
 */
#ifndef PUSCH_DEC_CUDA_H
#define PUSCH_DEC_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDC_ABI_VERSION 9

typedef enum {
  PDC_OK               = 0,
  PDC_ERR_INVALID_ARG  = 1, /* a contract the reference guards with srsran_assert was violated */
  PDC_ERR_NO_DEVICE    = 2,
  PDC_ERR_CUDA         = 3, /* see pdc_last_error() */
  PDC_ERR_CAPACITY     = 4, /* batch larger than the context was created for / no free ticket */
  PDC_ERR_NOT_READY    = 5, /* pdc_poll only */
  PDC_ERR_BAD_TICKET   = 6
} pdc_status;

/* CB CRC selection (crc_calculator_generic_impl.cpp:31-56). */
#define PDC_CRC_NONE 0 /* ldpc_decoder::decode(crc = nullptr): run max_iter iterations, no early stop */
#define PDC_CRC16 1
#define PDC_CRC24A 2
#define PDC_CRC24B 3

/* pdc_cb_desc.flags */
#define PDC_F_NEW_DATA 0x01u       /* first HARQ transmission: copy (not combine) the first lap (rate_dematch new_data) */
#define PDC_F_EARLY_STOP 0x02u     /* check the CB CRC after every iteration (pusch_codeblock_decoder use_early_stop) */
#define PDC_F_FORCE_DECODING 0x04u /* ldpc_decoder_factory_configuration::force_decoding (semantics as implemented) */
#define PDC_F_GENERIC 0x08u        /* arithmetic of ldpc_decoder_generic (scale = round(0.8f x)); default AVX2/AVX512 */
#define PDC_F_DECODE_ONLY 0x10u    /* ldpc_decoder::decode: llrs are the E = N_in de-matched soft bits, no HARQ buffer */
#define PDC_F_DEMATCH_ONLY 0x20u   /* ldpc_rate_dematcher::rate_dematch only (CB already decoded: combine, no decode) */
#define PDC_F_DECODED 0x40u        /* pdc_*_tbs only: the CB passed its CRC in an earlier submission and is not enqueued again
                                      (pusch_decoder_hw_impl.cpp:236-243): no LLRs are read; its bits, kept with the HARQ slot
                                      (harq_slot >= 0 required), take part in the transport-block assembly */

/* One code block. Mirrors codeblock_metadata (include/srsran/phy/upper/codeblock_metadata.h:41-100) plus
 * ldpc_decoder::configuration (ldpc_decoder.h:44-55) and hw_pusch_decoder_configuration
 * (hw_accelerator_pusch_dec.h:39-72). */
typedef struct {
  uint8_t  bg;         /* 1 | 2 */
  uint8_t  Qm;         /* 1, 2, 4, 6, 8 */
  uint8_t  rv;         /* 0..3 */
  uint8_t  crc_type;   /* PDC_CRC_* */
  uint8_t  max_iter;   /* > 0 */
  uint8_t  flags;      /* PDC_F_* */
  uint16_t Z;          /* lifting size */
  uint32_t E;          /* rate-matched length in LLRs (decode-only: number of input soft bits, multiple of Z) */
  uint32_t Nref;       /* limited-buffer rate matching length, 0 = none */
  uint32_t nof_filler; /* filler bits F */
  int32_t  harq_slot;  /* >= 0: soft buffer lives in the device arena slot (external HARQ); -1: exchanged with host */
  uint64_t llr_offset; /* byte offset of this CB's LLRs in the llrs buffer */
  uint64_t soft_offset;/* byte offset of this CB's N-entry soft buffer in the soft buffers (harq_slot < 0 only) */
  uint64_t out_offset; /* byte offset of this CB's packed hard bits (ceil(K/8) bytes, MSB first) in the bits buffer;
                          must be a multiple of 4; the bytes up to the next multiple of 4 belong to the CB as padding */
} pdc_cb_desc;

/* Per-CB outcome: hw_pusch_decoder_outputs (hw_accelerator_pusch_dec.h:75-80) / optional<unsigned> of decode(). */
typedef struct {
  int32_t  iterations; /* 1-based iteration at which the CRC matched; max_iter if it matched only at the end (no
                          early stop); -1 = reference's std::nullopt (CRC failed, or crc_type NONE) */
  uint8_t  crc_pass;   /* 1 iff iterations > 0 */
  uint8_t  status;     /* 0 ok; 1 = descriptor rejected (PDC_ERR_INVALID_ARG for this CB only) */
  uint16_t nof_layers; /* base-graph check nodes actually processed (ldpc_decoder_impl.cpp:115), 0 if not decoded */
} pdc_cb_result;

/* One transport block = C consecutive code blocks of the same submission (pusch_decoder_impl.cpp:396-505). All of them
 * share bg, Z, nof_filler and crc_type; C == 1: crc_type CRC16 | CRC24A and tbs_bits = K - F - crc bits; C > 1: CRC24B
 * and (C - 1) D + 24 < tbs_bits + 24 <= C D with D = K - F - 24 data bits per code block: the payload and its CRC24A fill
 * the code blocks in order, what is left of the last one is zero padding (TS 38.212 5.2.2 as
 * ldpc_segmenter_rx_impl.cpp:52-106 derives it; pusch_decoder_impl.cpp:485-487). */
typedef struct {
  uint32_t first_cb;  /* index of the TB's first code block in descs[] */
  uint32_t nof_cbs;   /* C >= 1 */
  uint32_t tbs_bits;  /* transport block size in bits (multiple of 8), TB CRC not included */
  uint32_t reserved;  /* 0 */
  uint64_t tb_offset; /* byte offset of the tbs_bits / 8 payload bytes in the TB buffer; multiple of 4 */
} pdc_tb_desc;

/* pusch_decoder_result (pusch_decoder_result.h) for one TB. A CB counts as passed if its CRC matched in this submission
 * or it carries PDC_F_DECODED / PDC_F_DEMATCH_ONLY. */
typedef struct {
  uint8_t  tb_crc_ok;  /* C == 1: the CB's CRC; C > 1: every CB passed AND CRC24A(payload) equals the 24 bits that follow it */
  uint8_t  written;    /* 1 iff the payload bytes were written: every CB passed (on a TB CRC mismatch they ARE written and
                          the caller resets its CB CRC flags: pusch_decoder_impl.cpp:431-437) */
  uint16_t nof_cbs_ok; /* code blocks of the TB that count as passed */
} pdc_tb_result;

typedef struct pdc_ctx pdc_ctx;
typedef int64_t        pdc_ticket;

typedef struct {
  int      device;        /* CUDA device ordinal */
  uint32_t max_batch_cbs; /* largest number of CBs in one submit */
  uint64_t max_batch_llrs;/* largest total number of LLR bytes in one submit */
  uint32_t harq_slots;    /* device HARQ arena capacity in CBs (each slot = 25344 B = 66 * 384) */
  uint32_t nof_tickets;   /* submits that may be in flight at once (each owns pinned + device staging) */
} pdc_config;

/* --- lifetime ---------------------------------------------------------------------------------------------------- */
/* Diagnostics: with PDC_TRACE=1 in the environment at pdc_create, every pdc_submit* records timing events around its
 * stages and pdc_destroy prints the averages on stderr (descriptor upload, kernels, device-to-host copies, and the time
 * from the end of the submit call until the host saw the completion). Off by default: no event is created. */
pdc_status  pdc_create(const pdc_config* cfg, pdc_ctx** out);
void        pdc_destroy(pdc_ctx* ctx);
const char* pdc_last_error(const pdc_ctx* ctx); /* copy owned by the calling thread, valid until its next call; "" if none */
/* Zero-copy submission (off by default). Off: pdc_submit* copies the LLRs into the ticket's pinned staging before it
 * returns, so the caller's memory may be reused at once. On: LLRs that lie in page-locked host memory (cudaHostAlloc /
 * cudaHostRegister, pdc_host_alloc) are read by the DMA engine from where they lie AFTER the call has returned: the caller
 * must leave them untouched until pdc_poll returns PDC_OK or pdc_fetch* has returned. Pageable memory is always staged. */
pdc_status  pdc_set_zero_copy(pdc_ctx* ctx, int enable);
/* Page-locked host memory for zero-copy submissions (cudaHostAlloc / cudaFreeHost behind the C ABI, for hosts that do not
 * link the CUDA runtime themselves). */
void*       pdc_host_alloc(size_t bytes);
void        pdc_host_free(void* p);
int         pdc_abi_version(void);

/* --- host-buffer path (what the srsRAN adapters call): H2D copy, kernel, D2H copy on one stream ---------------- */
/* Queues n CBs. llrs: host buffer holding every CB's LLRs (desc.llr_offset). soft_in: host soft buffers for CBs with
 * harq_slot < 0 (NULL allowed when every CB has PDC_F_NEW_DATA|harq_slot>=0|DECODE_ONLY). Copies into pinned staging,
 * so the caller's memory may be reused as soon as the call returns — unless pdc_set_zero_copy is on and llrs is page-locked
 * (see there). Returns PDC_ERR_CAPACITY when every ticket is in flight: back-pressure, not a failure; retry. */
pdc_status pdc_submit(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const int8_t* llrs, size_t llr_bytes,
                      const int8_t* soft_in, size_t soft_bytes, pdc_ticket* ticket);
/* Staged submission: a staging arena is page-locked host memory mirrored 1:1 by device memory, with a copy stream of its own.
 * Its users (the accelerator objects of the reference's pools, through the adapters' batcher) write a transport block's LLRs
 * into a region of the host side and push the region as soon as it is complete; the submission that later decodes it
 * (pdc_submit_staged: desc.llr_offset relative to the arena) only waits for what has been pushed — the host-to-device copy
 * of a batch overlaps the time its other members are still being filled. The bbdev counterpart is the mbuf pool an
 * accelerator enqueues from (hw_accelerator_factories.h:33-51). A region must not be rewritten before the submission that
 * used it has completed. */
typedef struct pdc_stage pdc_stage;
pdc_status pdc_stage_open(pdc_ctx* ctx, size_t bytes, pdc_stage** out);
void       pdc_stage_close(pdc_stage* stage);
int8_t*    pdc_stage_host(pdc_stage* stage);                                /* host side of the arena */
pdc_status pdc_stage_push(pdc_stage* stage, size_t offset, size_t bytes);   /* asynchronous host-to-device copy of a region */
pdc_status pdc_submit_staged(pdc_ctx* ctx, pdc_stage* stage, const pdc_cb_desc* descs, uint32_t n, const int8_t* soft_in,
                             size_t soft_bytes, pdc_ticket* ticket);        /* then pdc_poll / pdc_fetch as usual */
/* PDC_OK when the batch has completed, PDC_ERR_NOT_READY otherwise. */
pdc_status pdc_poll(pdc_ctx* ctx, pdc_ticket ticket);
/* Blocks until done, then scatters the packed bits to bits + desc.out_offset, the results to res[0..n) and, for CBs
 * with harq_slot < 0 (and not DECODE_ONLY), the updated soft buffers to soft_out + desc.soft_offset (NULL = skip). */
pdc_status pdc_fetch(pdc_ctx* ctx, pdc_ticket ticket, uint8_t* bits, size_t bits_bytes, pdc_cb_result* res,
                     int8_t* soft_out, size_t soft_bytes);

/* --- transport-block path: code blocks decoded AND assembled on the device, one payload per TB comes back ---------------
 * As pdc_submit, plus ntb transport blocks over those code blocks. Code blocks with harq_slot >= 0 keep their decoded bits
 * with the HARQ slot, so that a later transmission that re-sends only the failed ones (the others flagged PDC_F_DECODED)
 * still assembles the whole TB on the device. desc.out_offset is used only by CBs with harq_slot < 0. */
pdc_status pdc_submit_tbs(pdc_ctx* ctx, const pdc_cb_desc* descs, uint32_t n, const pdc_tb_desc* tbs, uint32_t ntb,
                          const int8_t* llrs, size_t llr_bytes, const int8_t* soft_in, size_t soft_bytes,
                          pdc_ticket* ticket);
/* Blocks until done; TB payloads to tb_bytes + tb.tb_offset (only TBs with written == 1), TB results to tb_res[0..ntb),
 * per-CB results (iteration statistics) to cb_res[0..n) if not NULL, soft buffers as pdc_fetch. No per-CB bits travel. */
pdc_status pdc_fetch_tbs(pdc_ctx* ctx, pdc_ticket ticket, uint8_t* tb_bytes, size_t tb_bytes_len, pdc_tb_result* tb_res,
                         pdc_cb_result* cb_res, int8_t* soft_out, size_t soft_bytes);

/* --- device-resident path (inputs and outputs already in HBM; used for kernel-only measurements and by callers that
 *     produce LLRs on the GPU). All pointers are device pointers; stream is a cudaStream_t (NULL = default). ------- */
pdc_status pdc_decode_device(pdc_ctx* ctx, const pdc_cb_desc* descs_host, uint32_t n, const int8_t* d_llrs,
                             int8_t* d_soft, uint8_t* d_bits, pdc_cb_result* d_res, void* stream);

/* Device-resident transport-block variant: d_bits may be NULL when every CB has harq_slot >= 0. */
pdc_status pdc_decode_device_tbs(pdc_ctx* ctx, const pdc_cb_desc* descs_host, uint32_t n, const pdc_tb_desc* tbs_host,
                                 uint32_t ntb, const int8_t* d_llrs, int8_t* d_soft, uint8_t* d_bits, pdc_cb_result* d_res,
                                 uint8_t* d_tb_bytes, pdc_tb_result* d_tb_res, void* stream);

/* Allocation-free variant of pdc_decode_device_tbs. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same arguments, same results. The call allocates nothing on the device or through any library: the device
 * scratch of the path (device descriptors, launch order, every code block's share of the TB CRC between
 * pdc_tb_assemble_kernel and pdc_tb_finish_kernel) lives in one file-scope `static __device__` array of 112 KiB that every
 * GPU loading the library holds until exit; its address is resolved once per device with cudaGetSymbolAddress.
 * Limit: n <= PDC_STATIC_MAX_CBS and ntb <= PDC_STATIC_MAX_CBS, else PDC_ERR_CAPACITY (pdc_last_error says why).
 * Every call on a device shares that array (and keeps the previous call's contents in it; what must start at zero is
 * cleared on `stream` by the call): calls are serialised on the host, but two calls whose GPU work OVERLAPS (different
 * streams without an ordering between them) race on the scratch - issue them on one stream or order the streams. */
#define PDC_STATIC_MAX_CBS 1024
pdc_status pdc_decode_device_tbs_static(pdc_ctx* ctx, const pdc_cb_desc* descs_host, uint32_t n, const pdc_tb_desc* tbs_host,
                                        uint32_t ntb, const int8_t* d_llrs, int8_t* d_soft, uint8_t* d_bits,
                                        pdc_cb_result* d_res, uint8_t* d_tb_bytes, pdc_tb_result* d_tb_res, void* stream);

/* --- LLR producers on the device: soft demodulation + descrambling --------------------------------------
 * One segment = ONE call of demodulation_mapper::demodulate_soft (in pusch_demodulator_impl.cpp:370-395 the resource elements
 * of one OFDM symbol), optionally followed by revert_scrambling with the TS 38.211 5.2.1 sequence of c_init = rnti 2^15 + n_id
 * (:261-262). The reference's result depends on a symbol's index within the call (AVX512 batches, then AVX2 batches, then
 * single symbols, each with its own rounding); the kernel reproduces the AVX512-host build bit for bit, so segment boundaries
 * must be the reference's call boundaries. The int8 LLRs land where pdc_decode_device reads them: no PCIe hop in between. */
#define PDC_MOD_PI2_BPSK 0
#define PDC_MOD_BPSK 1
#define PDC_MOD_QPSK 2
#define PDC_MOD_QAM16 4
#define PDC_MOD_QAM64 6
#define PDC_MOD_QAM256 8
#define PDC_DEMOD_F_DESCRAMBLE 0x01u /* multiply LLR i of the segment by 1 - 2 c(scramble_offset + i) */

typedef struct {
  uint64_t sym_offset;      /* index of the segment's first symbol in symbols[] (complex) and noise_vars[] */
  uint64_t llr_offset;      /* byte offset of its first LLR in llrs[]; nof_symbols * bits-per-symbol bytes are written */
  uint32_t nof_symbols;
  uint32_t c_init;          /* scrambling sequence initialiser (31 bits) */
  uint32_t scramble_offset; /* sequence bits consumed before this segment (the LLRs of the codeword that precede it) */
  uint8_t  mod;             /* PDC_MOD_* */
  uint8_t  flags;           /* PDC_DEMOD_F_* */
  uint16_t reserved;        /* 0 */
} pdc_demod_seg;

/* Device-resident: d_symbols = interleaved (re, im) floats (cf_t), d_noise_vars = one float per symbol (the equaliser's
 * post-equalisation noise variances), d_llrs = int8 output; asynchronous on stream (a cudaStream_t, NULL = default). */
pdc_status pdc_demodulate_device(pdc_ctx* ctx, const pdc_demod_seg* segs_host, uint32_t n, const float* d_symbols,
                                 const float* d_noise_vars, int8_t* d_llrs, void* stream);
/* Host buffers (what a demodulation_mapper adapter calls): copies in, runs, copies out, returns when the LLRs are in llrs.
 * nof_symbols / llr_bytes: sizes of the host arrays (every segment must lie inside). */
pdc_status pdc_demodulate(pdc_ctx* ctx, const pdc_demod_seg* segs, uint32_t n, const float* symbols, const float* noise_vars,
                          size_t nof_symbols, int8_t* llrs, size_t llr_bytes);

/* --- UL-SCH demultiplexer: ulsch_demultiplex (ulsch_demultiplex.h:36-103, ulsch_demultiplex_impl.cpp) ---------
 * Splits the descrambled soft bits of one PUSCH codeword into the UL-SCH data stream (what the code-block chain consumes) and the
 * HARQ-ACK / CSI Part 1 / CSI Part 2 streams (TS 38.212 6.2.7), with the reference's handling of the reserved REs (1- or 2-bit
 * HARQ-ACK punctures them: those soft bits reach the other streams as zeros) and of the x / y placeholder bits (their
 * scrambling is undone: needs c_init). One descriptor per codeword carries ulsch_demultiplex::configuration; the input is the
 * whole allocation in demodulation order (OFDM symbol by symbol, the output of pdc_demodulate* with PDC_DEMOD_F_DESCRAMBLE,
 * scramble_offset == position in the codeword). A PUSCH without UCI needs no call: its soft bits are the UL-SCH stream. */
#define PDC_DEMUX_F_CSI2_AFTER_CSI1 0x01 /* CSI Part 2 REs are allocated from the OFDM symbol in which CSI Part 1 ends, as when
                                            set_csi_part2() is called from the CSI Part 1 feedback (pusch_processor_impl.cpp:73-98);
                                            without the flag: from the first symbol (set right after demultiplex(), TS 38.212) */
typedef struct {
  uint8_t  Qm;                          /* 1, 2, 4, 6, 8 */
  uint8_t  nof_layers;                  /* 1..4 */
  uint8_t  start_symbol_index;          /* first OFDM symbol of the allocation */
  uint8_t  nof_symbols;                 /* start_symbol_index + nof_symbols <= 14 */
  uint8_t  dmrs_type;                   /* 1 | 2 */
  uint8_t  nof_cdm_groups_without_data; /* 1..2 (type 1), 1..3 (type 2) */
  uint16_t dmrs_symbol_mask;            /* bit l: OFDM symbol l of the slot carries DM-RS */
  uint16_t nof_prb;                     /* 1..275 */
  uint8_t  flags;                       /* PDC_DEMUX_F_* */
  uint8_t  reserved;                    /* 0 */
  uint32_t nof_harq_ack_rvd;            /* soft bits reserved for HARQ-ACK (0 unless nof_harq_ack_bits <= 2) */
  uint32_t nof_harq_ack_bits, nof_enc_harq_ack_bits;
  uint32_t nof_csi_part1_bits, nof_enc_csi_part1_bits;
  uint32_t nof_csi_part2_bits, nof_enc_csi_part2_bits;
  uint32_t c_init;                      /* scrambling initial state of the codeword */
  uint32_t reserved2;                   /* 0 */
  uint64_t in_offset;                   /* byte offset of the codeword's first soft bit in the input buffer */
  uint64_t sch_offset;                  /* byte offsets of the four streams in the output buffer */
  uint64_t harq_ack_offset;
  uint64_t csi_part1_offset;
  uint64_t csi_part2_offset;
} pdc_ulsch_demux_desc;

/* Soft bits of the codeword and of each stream: lens[0] input, [1] UL-SCH, [2] HARQ-ACK, [3] CSI Part 1, [4] CSI Part 2.
 * PDC_ERR_INVALID_ARG if the configuration is inconsistent (the reference asserts: UCI that does not fit the allocation,
 * encoded lengths that are not whole REs, ...). Host-only. */
pdc_status pdc_ulsch_demux_lengths(const pdc_ulsch_demux_desc* d, uint32_t lens[5]);
/* Device pointers, asynchronous on stream. Streams of different codewords / kinds must not overlap in d_out. */
pdc_status pdc_ulsch_demultiplex_device(pdc_ctx* ctx, const pdc_ulsch_demux_desc* descs_host, uint32_t n, const int8_t* d_in,
                                        int8_t* d_out, void* stream);
/* Host buffers: copies in, runs, copies the streams out, returns when they are in `out`. */
pdc_status pdc_ulsch_demultiplex(pdc_ctx* ctx, const pdc_ulsch_demux_desc* descs, uint32_t n, const int8_t* in, size_t in_len,
                                 int8_t* out, size_t out_len);

/* --- PDSCH transport-block encoder (the downlink mirror) ----------------------------------------------------
 * pdsch_encoder::encode (pdsch_encoder.h:48-73, pdsch_encoder_impl.cpp:31-80): TB CRC attachment, segmentation, code-block
 * CRC, LDPC encoding, rate matching + bit interleaving, code-block concatenation. One descriptor per transport block carries
 * pdsch_encoder::configuration; the segmentation (C, Z, K', fillers, per-code-block rate-matched lengths: ldpc.h:128-214,
 * ldpc_segmenter_helpers.h:82-100) is derived inside. The codeword comes back as the reference delivers it: one bit per
 * byte, nof_ch_symbols * Qm bytes (nof_ch_symbols counts the resource elements of every layer, as in the reference). */
typedef struct {
  uint8_t  bg;             /* 1 | 2 */
  uint8_t  Qm;             /* 1, 2, 4, 6, 8 (modulation_scheme as bits per symbol) */
  uint8_t  rv;             /* 0..3 */
  uint8_t  nof_layers;     /* 1..4 (ignored when nof_ch_symbols == 0) */
  uint32_t tbs_bits;       /* transport block size in bits (multiple of 8) */
  uint32_t nof_ch_symbols; /* channel symbols the transport block is mapped to, all layers together (multiple of nof_layers);
                              0 = the rate-matched lengths below are given explicitly */
  uint32_t Nref;           /* limited-buffer rate matching length, 0 = none */
  uint64_t tb_offset;      /* byte offset of the packed transport block (MSB first) in the input buffer */
  uint64_t cw_offset;      /* byte offset of the codeword in the output buffer */
  /* Explicit rate-matched lengths, as hal::hw_pdsch_encoder_configuration carries them (hw_accelerator_pdsch_enc.h:44-51:
   * that structure has no layer count, so E_r cannot be re-derived). Used only when nof_ch_symbols == 0, all 0 otherwise:
   * the first nof_short_segments code blocks get cw_length_a bits, the others cw_length_b (multiples of Qm). */
  uint32_t cw_length_a;
  uint32_t cw_length_b;
  uint32_t nof_short_segments;
  uint32_t reserved;       /* 0 */
} pdc_pdsch_tb_desc;

pdc_status pdc_pdsch_encode(pdc_ctx* ctx, const pdc_pdsch_tb_desc* tbs, uint32_t n, const uint8_t* tb_bytes, size_t tb_len,
                            uint8_t* codewords, size_t cw_len); /* host buffers, returns when the codewords are there */
pdc_status pdc_pdsch_encode_device(pdc_ctx* ctx, const pdc_pdsch_tb_desc* tbs_host, uint32_t n, const uint8_t* d_tb_bytes,
                                   uint8_t* d_codewords, void* stream); /* device pointers, asynchronous on stream */

/* The segmentation pdc_pdsch_encode will use for a transport block (TS 38.212 5.2.2 as ldpc.h:128-214 and
 * ldpc_segmenter_tx_impl.cpp derive it): number of code blocks C, lifting size Z, bits per code block before fillers K'
 * (information + CRCs), filler bits K - K'. Any output pointer may be NULL. Host-only, no device work. */
pdc_status pdc_pdsch_segmentation(int bg, uint32_t tbs_bits, uint32_t* nof_cbs, uint32_t* lifting_size, uint32_t* cb_info_bits,
                                  uint32_t* nof_filler_bits);

/* --- HARQ arena (ext_harq_buffer_context_repository.h:36-107 / free_harq_context_entry) --------------------------- */
pdc_status pdc_harq_free(pdc_ctx* ctx, int32_t slot);                               /* zero-fills the slot */
/* The slot's content is no longer needed (hw_accelerator_pusch_dec::free_harq_context_entry, hw_accelerator_pusch_dec.h:107-110):
 * from now on it READS AS ZEROS. Host-side bookkeeping only, no CUDA call: a first transmission rewrites the whole buffer
 * anyway; any other operation on a released slot has it cleared on its stream first. With it, first transmissions under
 * limited-buffer rate matching (Nref < N: the reference leaves a stretch of the buffer untouched, ldpc_rate_dematcher_impl.cpp
 * :198-200) find that stretch empty and take the four-row kernel. */
pdc_status pdc_harq_release(pdc_ctx* ctx, int32_t slot);
pdc_status pdc_harq_read(pdc_ctx* ctx, int32_t slot, int8_t* host_dst, size_t n);   /* test / debug access */
pdc_status pdc_harq_write(pdc_ctx* ctx, int32_t slot, const int8_t* host_src, size_t n);

/* --- introspection ------------------------------------------------------------------------------------------------ */
uint64_t pdc_kernel_launches(const pdc_ctx* ctx); /* CUDA kernels launched by this context so far */
/* Measures the device's instruction issue rate (thread-level instructions per second) for the instruction classes
 * the decoder uses; mode: 0 LOP3, 1 IMAD, 2 LOP3+IMAD, 3 HMNMX2+HFMA2, 4 VIMNMX.S16x2+VIADD.16x2. This is the
 * roofline denominator reported by bench.py (the path is integer-ALU bound). */
pdc_status pdc_probe_issue_rate(pdc_ctx* ctx, int mode, double* thread_inst_per_s);
/* Geometry helpers mirroring ldpc.h: K = Kb*Z, N = 66Z | 50Z; 0 if (bg, Z) is invalid. */
uint32_t pdc_cb_message_bits(int bg, int Z);
uint32_t pdc_cb_buffer_len(int bg, int Z);
/* Upper bound on the base-graph layers the decoder will process for this descriptor (sizes shared memory). */
uint32_t pdc_cb_max_layers(const pdc_cb_desc* d);

#ifdef __cplusplus
}
#endif
#endif /* PUSCH_DEC_CUDA_H */
