// nt4_swar.h -- branch-free ASCII -> 2-bit code + validity, four bytes at a time.
//
// Restates constants::seq_nt4_table (reference include/constants.hpp:12-21):
//   A,a -> 0   C,c -> 1   G,g -> 2   T,t,U,u -> 3   every other byte -> 4 ("break").
// The usual (c >> 1) & 3 trick gives A0 C1 G3 T2, which is NOT this order; here
//   code = ((c >> 1) ^ (c >> 2)) & 3   ->  A0 C1 G2 T3 (and U3),
// and validity is decided exactly (only the ten letters above are valid), so '@', 'E', 'W' ...
// which alias a code are flagged bad. Plain C++ (no CUDA types) so that the same function is
// compiled for the device and checked exhaustively on the host by tests/test_host_side.py::test_nt4_swar_exhaustive.
#ifndef BIOLIB_B200_NT4_SWAR_H
#define BIOLIB_B200_NT4_SWAR_H

#include <stdint.h>

#if defined(__CUDACC__)
#define BLG_HD __host__ __device__ __forceinline__
#else
#define BLG_HD static inline

#endif

// x = four ASCII bytes, byte 0 (least significant) = first base.
// Returns codes in bits 0..7 (base q at bits [2q, 2q+2)) and bad flags in bits 8..15
// (base q at bit 8 + 2q; the odd bits stay 0 so both halves share one geometry).
BLG_HD uint32_t blg_pack4(uint32_t x)
{
    const uint32_t t = ((x >> 1) ^ (x >> 2)) & 0x03030303u;
    const uint32_t y = x & 0xDFDFDFDFu;                        // fold case
    const uint32_t e = (y ^ 0x40404040u) & 0xC8C8C8C8u;         // bits 7,6,3 must read 0,1,0
    const uint32_t g = (y >> 2) & ~(y >> 1);                    // bit0: b2 & !b1
    const uint32_t b4 = y >> 4;                                 // bit0: b4 (the T/U row)
    const uint32_t ok = (b4 & g) | (~b4 & y & ~g);              // bit0: tail bits spell A,C,G | T,U
    uint32_t bad = e | (~ok & 0x01010101u);                     // any set bit in a byte => bad
    const uint32_t u1 = bad | (bad >> 3);                       // fold bits 3,6,7 towards bit 0 (stays in-byte)
    bad = (u1 | (u1 >> 3) | (u1 >> 4)) & 0x01010101u;
    const uint32_t code8 = (t * 0x01041040u) >> 24;             // gather 4 x 2 bits
    const uint32_t bad8 = (bad * 0x01041040u) >> 24;            // same geometry, values 0/1
    return code8 | (bad8 << 8);
}

// The two halves of blg_pack4 on their own, for the fused kernels' fast path (most 16-base words hold valid letters only):
//   blg_codes4  the four 2-bit codes gathered into bits 0..7 (meaningful for valid bytes)
//   blg_bad4    a word whose byte q is non-zero iff base q is not one of AaCcGgTtUu -- exact, 9 logic operations: after folding
//               the case (y = x & 0xDF) the valid bytes are 0x41 0x43 0x47 (bit 4 clear) and 0x54 0x55 (bit 4 set), i.e.
//                 bits 7,6,5,3 read 0,1,0,0                         (y ^ 0x40) & 0xE8 == 0
//                 g = b2 & !b1 must equal b4                        A (001) C (011) G (111): g = 0;  T (100) U (101): g = 1
//                 b4 | b0                                           the A/C/G row needs bit 0
BLG_HD uint32_t blg_codes4(uint32_t x)
{
    const uint32_t t = ((x >> 1) ^ (x >> 2)) & 0x03030303u;
    return (t * 0x01041040u) >> 24;
}
BLG_HD uint32_t blg_bad4(uint32_t x)
{
    const uint32_t y = x & 0xDFDFDFDFu;
    const uint32_t s1 = y >> 1, s2 = y >> 2, s4 = y >> 4;
    const uint32_t frame = (y ^ 0x40404040u) & 0xE8E8E8E8u;
    const uint32_t g = s2 & ~s1;
    return frame | (((g ^ s4) | ~(s4 | y)) & 0x01010101u);
}

// Sixteen bases: four 32-bit words (w[0] = first four bases) -> 32 bits of codes, 32 bits of flags.
BLG_HD void blg_pack16(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t* codes, uint32_t* flags)
{
    const uint32_t p0 = blg_pack4(w0), p1 = blg_pack4(w1), p2 = blg_pack4(w2), p3 = blg_pack4(w3);
    *codes = (p0 & 0xFFu) | ((p1 & 0xFFu) << 8) | ((p2 & 0xFFu) << 16) | ((p3 & 0xFFu) << 24);
    *flags = (p0 >> 8) | ((p1 >> 8) << 8) | ((p2 >> 8) << 16) | ((p3 >> 8) << 24);
}

// Same classification, different geometry for the second-generation kernel: 32 bits of codes (base q at bits
// [2q, 2q+2)) and 16 bad bits (base q at bit q).
BLG_HD void blg_pack16_bits(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t* codes, uint32_t* bad16)
{
    const uint32_t p0 = blg_pack4(w0), p1 = blg_pack4(w1), p2 = blg_pack4(w2), p3 = blg_pack4(w3);
    *codes = (p0 & 0xFFu) | ((p1 & 0xFFu) << 8) | ((p2 & 0xFFu) << 16) | ((p3 & 0xFFu) << 24);
    // bad flags of blg_pack4 sit at bits 8, 10, 12, 14: squeeze each group of four to bits 0..3
    const uint32_t f = ((p0 >> 8) & 0x55u) | (((p1 >> 8) & 0x55u) << 8) | (((p2 >> 8) & 0x55u) << 16) | (((p3 >> 8) & 0x55u) << 24);
    uint32_t x = (f | (f >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    x = (x | (x >> 8)) & 0x0000FFFFu;
    *bad16 = x;
}


#endif
