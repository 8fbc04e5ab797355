"""ctypes loader for the CPU warp emulator (tests/emu) -- TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EMU_SO = os.path.join(HERE, "emu", "libsw_emu.so")
CSRC = os.path.join(ROOT, "mgl_b200", "csrc")


def build_emu(force=False):
    srcs = [os.path.join(HERE, "emu", "sw_emu.cpp"), os.path.join(CSRC, "sw_plan.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")]
    if force or not os.path.exists(EMU_SO) or any(os.path.getmtime(d) > os.path.getmtime(EMU_SO) for d in deps):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-I" + CSRC, "-o", EMU_SO] + srcs)
    return EMU_SO


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(build_emu())
        lib.emu_align_batch.restype = C.c_int
        lib.emu_align_batch.argtypes = [C.c_int] + [C.c_void_p] * 6 + [C.c_int] * 5 + [C.c_void_p] * 3 + [C.c_int, C.c_void_p]
        _lib = lib
    return _lib


OPS = "MIDNS"


def pack_batch(pairs):
    """pairs: list of (ref bytes, alt bytes) -> arena, ref_off, ref_len, alt_off, alt_len"""
    n = len(pairs)
    ref_len = np.array([len(r) for r, _ in pairs], dtype=np.uint32)
    alt_len = np.array([len(a) for _, a in pairs], dtype=np.uint32)
    sizes = np.empty(2 * n, dtype=np.uint64)
    sizes[0::2] = ref_len
    sizes[1::2] = alt_len
    offs = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.uint64)
    arena = np.frombuffer(b"".join(r + a for r, a in pairs), dtype=np.uint8).copy()
    return arena, offs[0::2].copy(), ref_len, offs[1::2].copy(), alt_len


def emu_align(pairs, params, strategies, allow_s16=True, device_plan=False, clean_shortcut=True, through_state=False):
    """-> list of (cigar text, offset), plus stats dict.  device_plan: the chunk is planned by the device planner's
    stage bodies (mgl_b200/csrc/sw_plan_dev.h) instead of the host planner (sw_plan.cpp)."""
    lib = _load()
    lib.emu_use_device_plan(1 if device_plan else 0)
    lib.emu_use_clean_shortcut(1 if clean_shortcut else 0)
    lib.emu_use_through_state(1 if through_state else 0)
    n = len(pairs)
    arena, ro, rl, ao, al = pack_batch(pairs)
    st = np.asarray(strategies, dtype=np.int32)
    if st.ndim == 0:
        st = np.full(n, int(st), dtype=np.int32)
    stride = int((rl + al).max()) + 4
    offsets = np.zeros(n, dtype=np.int32)
    n_ops = np.zeros(n, dtype=np.uint32)
    ops = np.zeros(n * stride, dtype=np.uint32)
    stats = np.zeros(7, dtype=np.uint32)
    rc = lib.emu_align_batch(n, arena.ctypes.data, ro.ctypes.data, rl.ctypes.data, ao.ctypes.data, al.ctypes.data,
                             st.ctypes.data, params[0], params[1], params[2], params[3], 1 if allow_s16 else 0,
                             offsets.ctypes.data, n_ops.ctypes.data, ops.ctypes.data, stride, stats.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"emu_align_batch rc={rc}")
    ops = ops.reshape(n, stride)
    out = []
    for p in range(n):
        text = "".join(f"{int(o) >> 4}{OPS[int(o) & 15]}" for o in ops[p, : n_ops[p]])
        out.append((text, int(offsets[p])))
    return out, dict(launches=int(stats[0]), s16=int(stats[1]), s32=int(stats[2]), units=int(stats[3]), builds=int(stats[4]), hops=int(stats[5]), clean=int(stats[6]))


# ---- PairHMM lane code (mgl_b200/csrc/ph_lane.h) on the CPU
PH_EMU_SO = os.path.join(HERE, "emu", "libph_emu.so")
_ph_lib = None


def build_ph_emu(force=False):
    srcs = [os.path.join(HERE, "emu", "ph_emu.cpp"), os.path.join(CSRC, "ph_tables.cpp")]
    deps = srcs + [os.path.join(CSRC, "ph_lane.h"), os.path.join(CSRC, "ph_tables.h")]
    if force or not os.path.exists(PH_EMU_SO) or any(os.path.getmtime(d) > os.path.getmtime(PH_EMU_SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", PH_EMU_SO] + srcs)
    return PH_EMU_SO


def ph_emu(reads, haps, use_double=False, rpl=4):
    """log10 likelihoods [n_reads, n_haps] from the emulated warp (rows per lane = rpl)"""
    global _ph_lib
    from oracle import pairhmm_oracle as P
    if _ph_lib is None:
        _ph_lib = C.CDLL(build_ph_emu())
        _ph_lib.ph_emu.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        _ph_lib.ph_emu.restype = C.c_int
    lengths, rb, hb = P.pack(reads, haps)
    out = np.zeros((len(reads), len(haps)), dtype=np.float64)
    _ph_lib.ph_emu(lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, int(use_double), int(rpl), out.ctypes.data)
    return out


def compare_planners(pairs_lens, params, strategies, allow_s16=True):
    """pairs_lens: [(tl, ql)].  Host planner vs device planner stage bodies on one chunk -> (rc, host8, device8):
    launches, units, s16 pairs, s32 pairs, checkpoint / last-row / carry / scratch words."""
    lib = _load()
    n = len(pairs_lens)
    rl = np.array([t for t, _ in pairs_lens], dtype=np.uint32)
    al = np.array([q for _, q in pairs_lens], dtype=np.uint32)
    sizes = np.empty(2 * n, dtype=np.uint64)
    sizes[0::2], sizes[1::2] = rl, al
    offs = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.uint64)
    ro, ao = offs[0::2].copy(), offs[1::2].copy()
    st = np.asarray(strategies, dtype=np.int32)
    if st.ndim == 0:
        st = np.full(n, int(st), dtype=np.int32)
    out = np.zeros(16, dtype=np.uint64)
    lib.emu_compare_planners.restype = C.c_int
    lib.emu_compare_planners.argtypes = [C.c_int] + [C.c_void_p] * 5 + [C.c_int] * 5 + [C.c_void_p]
    rc = lib.emu_compare_planners(n, ro.ctypes.data, rl.ctypes.data, ao.ctypes.data, al.ctypes.data, st.ctypes.data,
                                  params[0], params[1], params[2], params[3], 1 if allow_s16 else 0, out.ctypes.data)
    return rc, [int(x) for x in out[:8]], [int(x) for x in out[8:]]
