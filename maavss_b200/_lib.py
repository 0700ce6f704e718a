"""ctypes binding of libmaavss_b200.so (the C ABI in include/maavss_b200.h).

There is NO CPU fallback: if the shared library is missing this raises, and every compute
entry point needs CUDA tensors on a B200.  Prototypes are read from the header itself, so
the header is the single source of truth for the ABI.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
# MAAVSS_LIB selects another build of the SAME library (the debug variants of build.py); there is still no fallback.
LIB_PATH = os.environ.get("MAAVSS_LIB") or os.path.join(HERE, "libmaavss_b200.so")
HEADER = os.path.join(HERE, "..", "include", "maavss_b200.h")

_lib = None
_protos = None


def parse_header(path=HEADER):
    """Return {name: (restype, [argtypes])} for every prototype in the header."""
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"^\s*#.*$", "", txt, flags=re.M)
    protos = {}
    for m in re.finditer(r"(const char\*|long long|int|void)\s+(maavss_\w+)\s*\(([^;{}]*?)\)\s*;", txt, flags=re.S):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        argtypes = []
        for a in [x.strip() for x in args.replace("\n", " ").split(",")]:
            if a in ("void", ""):
                continue
            t = a.rsplit(" ", 1)[0].strip() if not a.endswith("*") else a
            stars = a.count("*")
            base = a.replace("const", "").replace("*", " ").split()
            base = " ".join(base[:-1]) if len(base) > 1 else base[0]
            base = base.replace("unsigned char", "uchar")
            if stars >= 1:
                argtypes.append(ctypes.POINTER(ctypes.c_int) if (base == "int" and stars == 1) else ctypes.c_void_p)
            elif base == "int":
                argtypes.append(ctypes.c_int)
            elif base == "long long":
                argtypes.append(ctypes.c_longlong)
            elif base == "float":
                argtypes.append(ctypes.c_float)
            elif base == "double":
                argtypes.append(ctypes.c_double)
            else:
                raise ValueError(f"unhandled C type in header: {a!r} ({name})")
        restype = {"int": ctypes.c_int, "long long": ctypes.c_longlong, "void": None, "const char*": ctypes.c_char_p}[ret]
        protos[name] = (restype, argtypes)
    return protos


def lib():
    global _lib, _protos
    if _profiler is not None:
        return _profiler
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is not built.  Run `python -m maavss_b200.build` (or __graft_entry__.build()); "
                "maavss_b200 has no CPU / PyTorch fallback.")
        L = ctypes.CDLL(LIB_PATH)
        _protos = parse_header()
        for name, (res, args) in _protos.items():
            fn = getattr(L, name)   # AttributeError here == header/library mismatch
            fn.restype = res
            fn.argtypes = args
        if L.maavss_abi_version() != 1:
            raise RuntimeError("libmaavss_b200.so ABI version mismatch")
        _lib = L
    return _lib


class _Profiler:
    """Optional per-call device timing (CUDA events on the launching stream) used by bench.py's roofline leg."""
    SKIP = {"maavss_last_error", "maavss_abi_version", "maavss_device_sm_count", "maavss_launch_count", "maavss_bn_splits",
            "maavss_reduce_workspace_bytes", "maavss_stft_out_dims", "maavss_stft_plan_create", "maavss_stft_plan_destroy",
            "maavss_dft_plan_create", "maavss_dft_plan_destroy", "maavss_gemm_tf32x3_supported", "maavss_gemm_tf32x3_workspace_bytes",
            "maavss_conv_tc_supported", "maavss_conv_tc_workspace_bytes", "maavss_conv_wgrad_tc_supported",
            "maavss_conv_wgrad_tc_workspace_bytes", "maavss_conv_tc3_supported", "maavss_conv_tc3_col_elems",
            "maavss_conv_tc3_workspace_bytes", "maavss_channel_sum_static_max_channels", "maavss_bilstm_recurrent_supported", "maavss_conv_tc3_permute_weights", "maavss_peer_capacity"}

    def __init__(self, real):
        self.real, self.records = real, []

    def __getattr__(self, name):
        fn = getattr(self.real, name)
        if name in self.SKIP or not name.startswith("maavss_"):
            return fn
        import torch

        def timed(*a):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*a)
            e1.record()
            self.records.append((name, a, e0, e1))
            return rc
        return timed


_profiler = None


def start_profile():
    global _profiler
    _profiler = _Profiler(lib())
    return _profiler


def stop_profile():
    """-> [(entry point, args, milliseconds)] of every call since start_profile()."""
    global _profiler
    import torch
    torch.cuda.synchronize()
    out = [(n, a, e0.elapsed_time(e1)) for n, a, e0, e1 in _profiler.records]
    _profiler = None
    return out


def check(rc):
    if rc != 0:
        msg = lib().maavss_last_error()
        raise RuntimeError(f"libmaavss_b200 error {rc}: {msg.decode() if msg else '?'}")


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def dptr(t, dtype=None):
    """Device pointer of a contiguous CUDA tensor (fails loudly otherwise)."""
    import torch
    if not t.is_cuda:
        raise RuntimeError("maavss_b200 ops need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("maavss_b200 ops need contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected dtype {dtype}, got {t.dtype}")
    return ctypes.c_void_p(t.data_ptr())


def fptr(t):
    import torch
    return dptr(t, torch.float32)
