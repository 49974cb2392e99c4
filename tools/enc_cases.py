"""Transport-block cases for the PDSCH encoder: (bg, tbs_bits, Qm, nof_layers, nof_ch_symbols, rv, Nref);
nof_ch_symbols counts every layer (a multiple of nof_layers), the codeword has nof_ch_symbols * Qm bits.
Chosen to cover: one and many code blocks, both base graphs, CRC16 / CRC24A transport blocks, fillers, every modulation order
and redundancy version, limited-buffer rate matching, wrap-around of the circular buffer (low rate), rate-matched lengths that
differ between code blocks, and the 152-code-block transport block of BASELINE config 2 mirrored to the downlink."""

CASES = [
    (2, 96, 2, 1, 150, 0, 0),
    (2, 320, 2, 1, 300, 1, 0),
    (2, 1000, 4, 1, 400, 2, 0),
    (2, 3824, 2, 2, 4200, 3, 0),        # largest CRC16 transport block, BG2
    (2, 3832, 6, 1, 1000, 0, 0),        # CRC24A, one code block
    (2, 7656, 2, 1, 4001, 0, 0),        # 3 code blocks, fillers, unequal rate-matched lengths
    (2, 7656, 4, 2, 3002, 2, 4000),     # limited buffer
    (1, 8424, 8, 1, 1200, 0, 0),        # one BG1 code block exactly (K' = 8448)
    (1, 8432, 8, 2, 1402, 0, 0),         # just over: 2 code blocks
    (1, 30000, 6, 2, 6002, 1, 0),
    (1, 30000, 6, 2, 18000, 3, 20000),   # low rate + limited buffer: wraps
    (1, 100000, 8, 4, 13204, 0, 0),
    (1, 100000, 4, 3, 105000, 2, 0),     # rate ~ 0.24: every parity row
    (1, 1277968, 8, 4, 170352, 0, 0),    # the largest transport block the reference's TX segmenter admits (tbs + 24 <= 1277992,
                                        # ldpc_segmenter_tx_impl.cpp:38-47): config 2 mirrored, 152 code blocks, Z = 384
]
