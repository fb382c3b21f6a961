"""
ctypes binding of libnichecompass_b200.so (the C ABI declared in include/nichecompass_b200.h).

There is NO fallback: if the library cannot be loaded, or a call returns a negative status,
an exception is raised.  Tensors are passed as raw device pointers; the CUDA stream is torch's
current stream.
"""
import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnichecompass_b200.so")

NC_F32, NC_BF16 = 0, 1

_T = {"p": ctypes.c_void_p, "i": ctypes.c_int32, "l": ctypes.c_int64, "f": ctypes.c_float,
      "d": ctypes.c_double, "s": ctypes.c_void_p}

HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "nichecompass_b200.h")


def _parse_header(path: str) -> dict:
    """name -> argument kinds in header order (p pointer, i int32, l int64, f float, d double,
    s stream).  The header is the single source of truth for the binding."""
    import re
    src = re.sub(r"/\*.*?\*/", "", open(path).read(), flags=re.S)
    out = {}
    for m in re.finditer(r"\b(int|int64_t|const char\*)\s+(nc_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        name, args = m.group(2), m.group(3).strip()
        kinds = ""
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                if "*" in a:
                    kinds += "s" if a.endswith("stream") else "p"
                elif a.startswith(("int64_t", "size_t")):
                    kinds += "l"
                elif a.startswith("float"):
                    kinds += "f"
                elif a.startswith("double"):
                    kinds += "d"
                else:
                    kinds += "i"
        out[name] = kinds
    return out


# entry points that return a value instead of a status code (bound separately in ``_load``)
VALUE_FUNCS = ("nc_last_error", "nc_version", "nc_launch_count", "nc_launch_count_add", "nc_gatv2_bwd_ws_bytes")
# synchronous host-side entry points (no stream argument; called through ``lib()`` directly, not ``call``)
HOST_FUNCS = ("nc_host_alloc", "nc_host_free", "nc_peer_alloc", "nc_peer_free", "nc_peer_get_handle", "nc_peer_open",
              "nc_peer_close")
SIGNATURES = {k: v for k, v in _parse_header(HEADER_PATH).items() if k not in VALUE_FUNCS}

_lib = None


class NicheCompassCudaError(RuntimeError):
    pass


def _load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        # build in-tree when a toolchain is present (builder container / first import on a box)
        try:
            from . import build as _build
            _build.build()
        except Exception as e:  # noqa: BLE001
            raise NicheCompassCudaError(
                f"{LIB_PATH} is missing and could not be built ({e}); nichecompass_b200 has no "
                "CPU fallback -- run `python -m nichecompass_b200.build`.") from e
    lib = ctypes.CDLL(LIB_PATH)
    lib.nc_last_error.restype = ctypes.c_char_p
    lib.nc_version.restype = ctypes.c_int
    lib.nc_launch_count.restype = ctypes.c_int64
    lib.nc_launch_count_add.restype = ctypes.c_int64
    lib.nc_launch_count_add.argtypes = [ctypes.c_int64]
    lib.nc_gatv2_bwd_ws_bytes.restype = ctypes.c_int64
    lib.nc_gatv2_bwd_ws_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32]
    for name, sig in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [_T[c] for c in sig]
    _lib = lib
    return lib


def lib():
    return _load()


def dtype_code(t: torch.dtype) -> int:
    if t == torch.float32:
        return NC_F32
    if t == torch.bfloat16:
        return NC_BF16
    raise TypeError(f"unsupported element type {t}")


def ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    if not t.is_cuda and not t.is_pinned():
        # pinned host memory is device-addressable under UVA (host-resident count store)
        raise NicheCompassCudaError("nichecompass_b200 kernels need CUDA tensors (no CPU fallback)")
    return ctypes.c_void_p(t.data_ptr())


class HostPtr:
    """Marks a CPU tensor passed to one of the few ``*_host`` arguments of the C ABI."""

    def __init__(self, t: torch.Tensor):
        assert not t.is_cuda and t.is_contiguous()
        self.t = t


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream() -> int:
    """cudaStream_t of torch's current stream on the current device (as an integer handle)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


_HOOK = None
_FN = {}
_Tensor = torch.Tensor


def set_call_hook(hook):
    """``hook.before(name, args)`` / ``hook.after(name, args)`` around every C-ABI call (bench.py's
    per-kernel CUDA-event timing); ``None`` removes it."""
    global _HOOK
    _HOOK = hook


def _fail(name: str, rc: int):
    msg = _lib.nc_last_error().decode()
    if rc in (-1, -2):
        raise ValueError(f"{name}: {msg}")
    raise NicheCompassCudaError(f"{name} failed ({rc}): {msg}")


def call(name: str, *args):
    """Call ``name`` with torch tensors (-> device pointers), None (-> NULL) or scalars; the
    stream argument is appended automatically.  This is the hot host path (~70 calls per
    training step): tensors are converted with one ``data_ptr()`` each, ctypes ``argtypes`` do
    the rest."""
    fn = _FN.get(name)
    if fn is None:
        fn = _FN[name] = getattr(_load(), name)
    conv = []
    for a in args:
        if isinstance(a, _Tensor):
            if not a.is_cuda and not a.is_pinned():
                # pinned host memory is device-addressable under UVA (host-resident count store)
                raise NicheCompassCudaError("nichecompass_b200 kernels need CUDA tensors (no CPU fallback)")
            conv.append(a.data_ptr())
        elif isinstance(a, HostPtr):
            conv.append(a.t.data_ptr())
        else:
            conv.append(a)
    if _raw_stream is not None:
        st = _raw_stream(torch.cuda.current_device())
    else:
        st = torch.cuda.current_stream().cuda_stream
    if _HOOK is not None:
        _HOOK.before(name, args)
        rc = fn(*conv, st)
        _HOOK.after(name, args)
    else:
        rc = fn(*conv, st)
    if rc != 0:
        _fail(name, rc)


def launch_count() -> int:
    return int(_load().nc_launch_count())
