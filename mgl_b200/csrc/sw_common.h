/*
 * sw_common.h -- shared definitions for the B200 Smith-Waterman path.
 *
 * Everything marked SW_HD compiles both as CUDA device code (the product) and
 * as plain host C++ (used only by the lock-step warp emulator under tests/emu,
 * which lets the fill/traceback logic be checked against the oracle on a
 * machine without a GPU).  The product never executes the host variants.
 *
 * Reference semantics restated here (paths relative to
 * /root/reference/src/main/native/mgl_sw/):
 *   overhang strategy bits          sw_common.h:22-25
 *   CIGAR state characters          sw_common.h:27-30
 *   score sign normalisation        com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman.cpp:52-55
 */
#ifndef MGL_B200_SW_COMMON_H
#define MGL_B200_SW_COMMON_H

#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define SW_HD __host__ __device__ __forceinline__
#define SW_D __device__ __forceinline__
#else
#define SW_HD inline
#define SW_D inline
#endif

namespace mglsw {

typedef uint32_t u32;
typedef uint64_t u64;

/* overhang strategies: sw_common.h:22-25 */
enum : int { OS_SOFTCLIP = 1, OS_INDEL = 2, OS_LEADING_INDEL = 4, OS_IGNORE = 8 };

/* packed CIGAR element = (length << 4) | op, BAM op numbering */
enum : u32 { OP_M = 0, OP_I = 1, OP_D = 2, OP_S = 4 };

/* which arithmetic a unit is filled with */
enum : int { KIND_S16X2 = 0, KIND_S32 = 1 };

/* scores after the reference's sign normalisation (...SmithWaterman.cpp:52-55) */
struct Scores {
    int match;      /* > 0 */
    int mismatch;   /* <= 0 */
    int go;         /* >= 0, gap open penalty magnitude   */
    int ge;         /* >= 0, gap extend penalty magnitude */
};

SW_HD Scores normalise_scores(int match, int mismatch, int gap_open, int gap_extend)
{
    Scores s;
    s.match = match > 0 ? match : -match;
    s.mismatch = mismatch < 0 ? mismatch : -mismatch;
    s.go = gap_open > 0 ? gap_open : -gap_open;
    s.ge = gap_extend > 0 ? gap_extend : -gap_extend;
    return s;
}

/* row-0 / column-0 boundary score, sw.cpp:29-35 and :47-49 */
SW_HD int boundary_score(int x, int indel, int go, int ge)
{
    return (indel && x >= 1) ? -go - (x - 1) * ge : 0;
}

/*
 * Constants of one fill launch (kernel parameter; lives in the constant bank).
 * For KIND_S16X2 every field holds the same 16-bit quantity in both halves.
 */
struct CellConsts {
    u32 neg_delta;  /* 32-bit multiplier: -(match - mismatch)                         */
    u32 neg_ge;     /* -gap_extend, replicated two's complement (per-half add-max)    */
    u32 match;      /* match score, replicated                                        */
    u32 one;        /* 1 as a RUNTIME value: x * one + y is then an integer multiply- */
                    /* add, which issues to the FMA pipe instead of the ALU pipe the  */
                    /* DPX instructions use                                           */
    u32 neg_go;     /* 0 - go: x + neg_go subtracts gap_open from both halves         */
    int bias;       /* value stored = true score + bias (0 for KIND_S32)              */
    int s_match, s_mismatch, s_go, s_ge;   /* plain scalars for boundary arithmetic and the traceback's recomputation */
};

/*
 * One lane group's work: one pair (KIND_S32) or two pairs of identical alt
 * length sharing every register as 16-bit halves (KIND_S16X2).
 */
struct UnitDesc {
    u64 ref_off[2];      /* byte offsets of the two refs in the sequence arena        */
    u64 alt_off[2];
    u64 ckpt_off;        /* word offset of the owning warp's checkpoint regions       */
    u64 reserved0;
    u64 carry_off;       /* uint2 index of this unit's two strip-carry columns        */
    u64 lastrow_off;     /* word offset of this unit's last-row capture               */
    u32 pair[2];         /* chunk-local pair ids; NO_PAIR: half unused                */
    u32 tl[2];           /* ref lengths                                               */
    u32 ql;              /* alt length (shared by both halves)                        */
    u32 tlmax;           /* rows the WARP iterates (max tl over its groups)           */
    u32 nstrips;         /* column strips (same for every group of the warp)          */
    u32 col_base;        /* columns per lane slot: col_base or col_base + 1           */
    u32 col_rem;         /* the first col_rem slots get col_base + 1                  */
    u32 rb_count;        /* row blocks (of MGLSW_RB steps) per strip                  */
    u32 carry_rows;      /* rows per carry column = own max tl + 1                    */
    u32 flags;           /* UNIT_* bits                                               */
    u32 group;           /* lane group index inside the warp                          */
    u32 regs;            /* K of the launch that fills this unit                      */
};
/* UNIT_INDEL: both pairs use the INDEL / LEADING_INDEL boundary (sw.cpp:29-40) */
enum : u32 { UNIT_INDEL = 1u, UNIT_ACTIVE = 4u, UNIT_G32 = 8u, UNIT_S32 = 16u };
static const u32 NO_PAIR = 0xffffffffu;

}  // namespace mglsw
#endif
