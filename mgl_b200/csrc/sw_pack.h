/*
 * sw_pack.h -- 2-bit transport form of a byte range of the arena.
 *
 * The aligner compares raw bytes (SW/sw.cpp:55: `ts[i-1] == qs[j-1]`), so the device arena always holds
 * the caller's bytes as they are.  What is packed is only the host-to-device TRANSPORT of a chunk's byte
 * range: when every byte of the range is one of A, C, G, T the planner thread writes four bases per byte
 * into pinned staging, a quarter of the bytes cross the host link, and `sw_unpack2_kernel` (sw_engine.cu)
 * writes the original bytes back into the arena.  A range holding any other byte (N, lower case, IUPAC
 * codes) is uploaded raw.  Results cannot depend on which way a range travelled.
 *
 * Code of a base = (byte >> 1) & 3: A 0, C 1, T 2, G 3.  Base k of the range sits in bits 2*(k%4).. of
 * packed byte k/4 (so a little-endian u32 holds 16 bases, base 0 in the low bits).
 */
#ifndef MGL_B200_SW_PACK_H
#define MGL_B200_SW_PACK_H

#include <cstddef>
#include <cstdint>

namespace mglsw {

/* bytes of the packed form of n bases, rounded up to whole u32 words */
inline size_t pack2_bytes(size_t n) { return ((n + 15) / 16) * 4; }

/*
 * Packs src[0, n) into dst (pack2_bytes(n) bytes; the unused high bits of the last word are zero).
 * Returns false, with dst partly written, as soon as a byte outside {A, C, G, T} is seen.
 * AVX2 when the CPU has it, a table loop otherwise; both produce the same bytes.
 */
bool pack2(const uint8_t *src, size_t n, uint8_t *dst);
bool pack2_scalar(const uint8_t *src, size_t n, uint8_t *dst);

/* the word the unpack kernel maps codes through: byte c of it is the base of code c */
constexpr uint32_t PACK2_BASES = 0x47544341u;   /* 'A' | 'C' << 8 | 'T' << 16 | 'G' << 24 */

}  // namespace mglsw
#endif
