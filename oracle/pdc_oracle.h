/* TEST INFRASTRUCTURE ONLY — declarations of the CPU oracle (see pdc_oracle.c for the status header). */
#ifndef PDC_ORACLE_H
#define PDC_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_FLAVOUR_GENERIC 0
#define ORC_FLAVOUR_AVX2 1
#define ORC_FLAVOUR_AVX512 2

#define ORC_CRC_NONE 0
#define ORC_CRC16 1
#define ORC_CRC24A 2
#define ORC_CRC24B 3

typedef struct {
  int             Kb, N_full, N_short, M, nedges;
  const uint8_t*  row_deg;
  const uint8_t*  edge_col;
  const uint16_t* edge_shift; /* row of the set index i_LS; reduce mod Z at use */
  int             row_start[47];
} orc_graph;

int      orc_lifting_set_index(int Z);
int      orc_graph_init(orc_graph* g, int bg, int Z);
uint32_t orc_crc_bits(int crc_id, const uint8_t* packed, int nbits);
int      orc_rate_dematch(int flavour, int8_t* buf, int N, const int8_t* in, int E, int new_data, int rv, int Qm,
                          int Nref, int nof_filler);
/* Returns the 1-based iteration count on CRC success, -1 for the reference's std::nullopt, -2 on contract violation. */
int orc_ldpc_decode(int flavour, int bg, int Z, int nof_filler, int max_iter, int crc_id, int force_decoding,
                    const int8_t* llrs, int n_llrs, uint8_t* out_packed);
/* pusch_codeblock_decoder::decode: dematch into buf, then decode; returns iterations or -1 (CRC failed). */
int orc_codeblock_decode(int flavour, int bg, int Z, int8_t* buf, const int8_t* in, int E, int new_data, int rv,
                         int Qm, int Nref, int nof_filler, int crc_id, int max_iter, int use_early_stop,
                         int force_decoding, uint8_t* out_packed);

/* pusch_decoder_impl::join_and_notify + concatenate_codeblocks (pusch_decoder_impl.cpp:396-505) for one transport block of
 * C code blocks with K message bits each (cb_packed: C buffers of cb_stride bytes, MSB first), nof_filler fillers, cb_crc_ok[C].
 * Writes tbs_bits / 8 bytes to tb_out only where the reference writes its transport_block span; returns tb_crc_ok and sets
 * *written (payload written) accordingly. */
int orc_tb_assemble(int C, int K, int nof_filler, const uint8_t* cb_packed, int cb_stride, const uint8_t* cb_crc_ok,
                    int tbs_bits, uint8_t* tb_out, int* written);

/* LLR producers, pdc_oracle_demod.c. mod: 0 pi/2-BPSK, 1 BPSK, 2 QPSK, 4 16QAM, 6 64QAM, 8 256QAM; symbols are
 * (re, im) float pairs; one call = one call of demodulation_mapper::demodulate_soft (the vector-batch boundaries follow it). */
void orc_demodulate_soft(int8_t* llrs, const float* symbols, const float* noise_vars, unsigned nof_symbols, int mod);
void orc_prg_bits(uint8_t* out, unsigned n, uint32_t c_init, unsigned offset);
void orc_descramble(int8_t* out, const int8_t* in, unsigned n, uint32_t c_init, unsigned offset);
extern unsigned orc_demod_fused_mask;

#ifdef __cplusplus
}
#endif
#endif
