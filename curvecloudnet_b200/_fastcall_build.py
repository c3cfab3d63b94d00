"""Builds `_ccn_cffi`, a CPython extension with one thin C wrapper per entry point of include/ccn_b200.h (cffi API mode).

Why: ctypes marshals every argument through libffi and Python-level converters - 2.8 us for a 9-argument call, 8.1 us for the
31 arguments of ccn_linear_planes, measured here - and the python executors make 180 (inference) to 690 (training step) calls
per step.  The generated wrappers take pointers as plain integers (`uintptr_t`, what `tensor.data_ptr()` returns) and cast
them back, so a call costs 0.4 - 0.9 us.  The extension links against libccn_b200.so (rpath $ORIGIN): same library, same
entry points, only the Python -> C hop differs; native.py falls back to ctypes when the extension is absent or was built
from another header (hash check).  Invoked by __graft_entry__.build()."""
import hashlib
import os
import re
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(HERE), "include", "ccn_b200.h")


def header_hash() -> str:
    return hashlib.sha1(open(HEADER, "rb").read()).hexdigest()


def generate():
    """-> (cdef text, C source text) from the header's declarations."""
    src = re.sub(r"/\*.*?\*/", " ", open(HEADER).read(), flags=re.S)
    cdef, wrap = ["const char* w_header_hash(void);"], []
    for ret, name, args in re.findall(r"\b(int|int64_t)\s+(ccn_\w+)\s*\(([^)]*)\)\s*;", src):
        params, casts = [], []
        for k, a in enumerate(a.strip() for a in args.split(",")):
            if a in ("", "void"):
                continue
            ctype = re.match(r"(.*?)(\w+)$", a).group(1).strip()          # the declaration minus the parameter name
            if "*" in ctype or "ccn_stream_t" in ctype or "ccn_draw_fn" in ctype:
                params.append("uintptr_t a%d" % k)
                casts.append("(%s)a%d" % (ctype, k))
            else:
                params.append("%s a%d" % (ctype, k))
                casts.append("a%d" % k)
        sig = ", ".join(params) or "void"
        cdef.append("%s w_%s(%s);" % (ret, name, sig))
        wrap.append("static %s w_%s(%s) { return %s(%s); }" % (ret, name, sig, name, ", ".join(casts)))
    csrc = ('#include <stdint.h>\n#include "ccn_b200.h"\nstatic const char* w_header_hash(void) { return "%s"; }\n'
            % header_hash()) + "\n".join(wrap) + "\n"
    return "\n".join(cdef), csrc


def build(verbose: bool = False) -> str:
    from cffi import FFI
    cdef, csrc = generate()
    ffi = FFI()
    ffi.cdef(cdef)
    ffi.set_source("_ccn_cffi", csrc, include_dirs=[os.path.dirname(HEADER)], libraries=["ccn_b200"], library_dirs=[HERE],
                   extra_link_args=["-Wl,-rpath,$ORIGIN"])
    tmp = os.path.join(HERE, "csrc", "build", "cffi")
    os.makedirs(tmp, exist_ok=True)
    built = ffi.compile(tmpdir=tmp, verbose=verbose)
    for stale in os.listdir(HERE):
        if stale.startswith("_ccn_cffi.") and stale.endswith(".so"):
            os.remove(os.path.join(HERE, stale))
    dst = os.path.join(HERE, os.path.basename(built))
    shutil.copy2(built, dst)
    return dst


if __name__ == "__main__":
    print(build(verbose=True))
