"""ctypes binding of libccn_b200.so (the C-ABI declared in include/ccn_b200.h).

This is the whole "thin torch-extension": torch supplies device memory and the current CUDA
stream, every piece of arithmetic happens in the library's sm_100a kernels.  There is NO
fallback: if the library is missing, or a tensor is not a CUDA tensor, calls raise.
"""
from __future__ import annotations

import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libccn_b200.so")

_c = {"p": ctypes.c_void_p, "q": ctypes.c_int64, "i": ctypes.c_int32, "f": ctypes.c_float, "d": ctypes.c_double}

def _parse_header():
    """Derive every entry point's ctypes signature from include/ccn_b200.h (single source of truth)."""
    path = os.path.join(os.path.dirname(_HERE), "include", "ccn_b200.h")
    src = re.sub(r"/\*.*?\*/", " ", open(path).read(), flags=re.S)
    sigs, rets = {}, {}
    for ret, name, args in re.findall(r"\b(int|int64_t)\s+(ccn_\w+)\s*\(([^)]*)\)\s*;", src):
        codes = ""
        for a in [a.strip() for a in args.split(",")]:
            if a in ("", "void"):
                continue
            if "*" in a or "ccn_stream_t" in a or "ccn_draw_fn" in a:
                codes += "p"
            elif "int64_t" in a:
                codes += "q"
            elif "int32_t" in a:
                codes += "i"
            elif "double" in a:
                codes += "d"
            elif "float" in a:
                codes += "f"
            else:
                raise RuntimeError("unparsed argument %r of %s" % (a, name))
        sigs[name] = codes
        rets[name] = ctypes.c_int64 if ret == "int64_t" else ctypes.c_int
    return sigs, rets


SIGNATURES, _RETURNS = _parse_header()

_lib = None


def load():
    """Load libccn_b200.so (built by __graft_entry__.build()); raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "curvecloudnet_b200: %s not built - run `python -c 'import __graft_entry__ as g; g.build()'`; "
            "there is no CPU fallback" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, sig in SIGNATURES.items():
        fn = getattr(lib, name)        # AttributeError here = the .so does not export a declared symbol
        fn.argtypes = [_c[ch] for ch in sig]
        fn.restype = _RETURNS[name]
    _lib = lib
    _load_fast()
    return lib


_fast = None                # the cffi extension's `lib` (thin C wrappers, see _fastcall_build.py) or None: ctypes only
BINDING = "ctypes"


def _load_fast():
    """Use the generated CPython extension for the Python -> C hop when it is present and was built from this header
    (CCN_BINDING=ctypes keeps ctypes).  Only the binding differs: both reach the same entry points of libccn_b200.so."""
    global _fast, BINDING
    if os.environ.get("CCN_BINDING", "") == "ctypes":
        return
    try:
        from . import _ccn_cffi
        from ._fastcall_build import header_hash
        mod_hash = _ccn_cffi.ffi.string(_ccn_cffi.lib.w_header_hash()).decode()
        if mod_hash != header_hash():
            import warnings
            warnings.warn("curvecloudnet_b200: _ccn_cffi was built from another include/ccn_b200.h - using ctypes "
                          "(run __graft_entry__.build())")
            return
        _fast, BINDING = _ccn_cffi.lib, "cffi"
    except ImportError:
        _fast = None


class CcnError(RuntimeError):
    pass


# ---- optional per-entry-point device timing (bench.py's roofline leg) -------------------------------
_prof_names = None
_prof_shapes = []          # (M, N, K) of every timed Linear call (tools/profile_forward.py)
_prof_events = []          # (name, start event, end event, algorithmic work of the call)

# algorithmic work of one call: FLOPs for the Linear kernels (2 M N K)
_MNK = {"ccn_linear_f32": lambda a: (a[7], a[8], a[9]), "ccn_linear_tc": lambda a: (a[7], a[8], a[9]),
        "ccn_linear_planes": lambda a: (a[24], a[25], a[3] + a[7] + a[11] + a[15]),
        "ccn_linear_wgrad": lambda a: (a[4], a[5], a[6]), "ccn_linear_wgrad_tc": lambda a: (a[6], a[7], a[8])}
_WORK = {name: (lambda a, f=f: 2.0 * f(a)[0] * f(a)[1] * f(a)[2]) for name, f in _MNK.items()}


def _is_buf(a):
    return a.__class__ is not int and a is not None


def _planes_bytes(a):
    """Algorithmic bytes of one ccn_linear_planes launch: every activation plane read once, the weights read once,
    every output (fp32 rows and / or bf16 planes) written once."""
    P = 2 if a[27] == 3 else 1
    M, N = a[24], a[25]
    kin = sum((a[i] + 7) // 8 * 8 for i in (3, 7, 11, 15))
    b = M * kin * 2 * P + N * (a[3] + a[7] + a[11] + a[15]) * 2 * P
    if _is_buf(a[18]):
        b += N * 4
    if _is_buf(a[19]):
        b += M * N * 4
    if _is_buf(a[21]):
        b += M * ((N + 15) // 16 * 16) * 2 * P
    return float(b)


# algorithmic bytes of one call (HBM roofline of the Linear kernels)
def _wgrad_bytes(a):
    """ccn_linear_wgrad_tc: dY planes and X planes read once, dW (fp32) accumulated once per row split (counted once)."""
    P = 2 if a[11] == 3 else 1
    M, N, K = a[6], a[7], a[8]
    return float(M * ((N + 7) // 8 * 8 + (K + 7) // 8 * 8) * 2 * P + N * K * 4)


_BYTES = {"ccn_linear_planes": _planes_bytes, "ccn_linear_wgrad_tc": _wgrad_bytes,
          "ccn_linear_wgrad": lambda a: 4.0 * (a[4] * (a[5] + a[6]) + a[5] * a[6]),
          "ccn_linear_f32": lambda a: 4.0 * (a[7] * a[9] + a[8] * a[9] + a[7] * a[8] + (a[8] if _is_buf(a[4]) else 0))}


def profile_begin(names, pad_cycles=0):
    global _prof_names, _prof_events, _prof_pad_cycles
    _prof_pad_cycles = int(pad_cycles)
    _prof_names = set(names) if names else None
    _prof_events = []
    del _prof_shapes[:]


def profile_end():
    """-> {name: {"ms": device time, "calls": n, "work": summed algorithmic work}} (synchronises)."""
    global _prof_names
    out = {}
    if _prof_events:
        torch.cuda.synchronize()
    for name, a, b, work, nbytes in _prof_events:
        d = out.setdefault(name, {"ms": 0.0, "calls": 0, "work": 0.0, "bytes": 0.0})
        d["ms"] += a.elapsed_time(b)
        d["calls"] += 1
        d["work"] += work
        d["bytes"] += nbytes
    _prof_names = None
    return out


_stream_handle = None      # set for the duration of a forward (ModelBase.forward) to skip the per-call lookup


def pin_stream(on: bool):
    """Cache torch's current CUDA stream handle for subsequent calls (the caller must not switch streams
    while pinned); pin_stream(False) returns to per-call lookup."""
    global _stream_handle
    _stream_handle = _current_raw_stream() if on else None


def _current_raw_stream() -> int:
    """torch's current CUDA stream handle of the current device (the C accessors: ~0.3 us instead of the ~15 us of
    torch.cuda.current_stream(), which matters in backward where no forward has pinned the stream)."""
    try:
        return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())
    except AttributeError:          # private accessors moved: fall back to the public API
        return torch.cuda.current_stream().cuda_stream


_fns = {}                   # name -> (foreign function, indices of the pointer arguments)


def _bind(name):
    lib = _lib if _lib is not None else load()
    sig = SIGNATURES[name]
    fn = getattr(_fast, "w_" + name, None) if _fast is not None else None
    ent = (fn if fn is not None else getattr(lib, name), tuple(i for i, ch in enumerate(sig[:-1]) if ch == "p"))
    _fns[name] = ent
    return ent


def call(name, *args):
    """Call a C-ABI entry point; torch tensors become device pointers, the current stream is appended."""
    ent = _fns.get(name) or _bind(name)
    conv = list(args)
    for i in ent[1]:
        a = conv[i]
        if a.__class__ is not int:
            if not a.is_cuda:
                raise CcnError("%s: tensor argument is not on a CUDA device (no CPU path exists)" % name)
            conv[i] = a.data_ptr()
    conv.append(_stream_handle if _stream_handle is not None else _current_raw_stream())
    if _prof_names is not None and name in _prof_names:
        return _timed_call(name, ent[0], conv, args)
    rc = ent[0](*conv)
    if rc != 0:
        _raise(name, rc)


def _raise(name, rc):
    e = CcnError("%s: invalid argument (code %d)" % (name, rc) if rc < 0 else "%s: CUDA error %d" % (name, rc))
    e.code = rc
    raise e


_prof_footprint = None      # list: argument footprint (bytes) of every timed call, parallel to _prof_events
_prof_pad_cycles = 0        # > 0: a spin of that many GPU cycles is queued before every timed launch, so that the
                            # launch is already in the queue when the GPU reaches the start event (the event pair then
                            # brackets the kernel, not the host's launch latency)


def _timed_call(name, fn, conv, args):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if _prof_pad_cycles:
        torch.cuda._sleep(_prof_pad_cycles)
    e0.record()
    rc = fn(*conv)
    e1.record()
    _prof_events.append((name, e0, e1, _WORK[name](args) if name in _WORK else 0.0,
                         _BYTES[name](args) if name in _BYTES else 0.0))
    if _prof_footprint is not None:
        # bytes of every tensor argument (inputs read + outputs written): the compulsory-traffic proxy of
        # tools/profile_forward.py's per-kernel GB/s table
        _prof_footprint.append(sum(a.numel() * a.element_size() for a in args if a.__class__ is not int and a.__class__ is not float))
    if name in _WORK:
        _prof_shapes.append(_MNK[name](args))
    if rc != 0:
        _raise(name, rc)


def kernel_launches() -> int:
    return int(load().ccn_kernel_launches())


def set_dependent_launch(on: int = -1) -> int:
    """ccn_set_dependent_launch: programmatic dependent launch of the library's kernels (1 / 0; -1 queries) -> previous."""
    return int(load().ccn_set_dependent_launch(int(on)))


def grid_num_cells(n2: int) -> int:
    return int(load().ccn_grid_num_cells(int(n2)))


def voxel_table_slots(n: int) -> int:
    return int(load().ccn_voxel_table_slots(int(n)))
