"""za_gpu -- thin ctypes binding over libza_gpu.so (the C ABI of include/za_gpu.h).

This is plumbing for tests and bench.py: the product is the shared library.  `BatchProcessor`
mirrors the reference host processor's call shape (prepareToPlay / processBlock,
src/JSFXJuceProcessor.cpp:3239, 3435) over n_streams instances.  There is no CPU fallback:
if the library is missing or no CUDA device exists, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent.parent
LIB_PATH = Path(os.environ.get("ZA_GPU_LIB", PKG / "libza_gpu.so"))   # override: A/B builds of the same ABI

ZA_F32, ZA_F64 = 0, 1
ZA_MEM_HOST, ZA_MEM_DEVICE = 0, 1


class ZaError(RuntimeError):
    pass


class PluginInfo(C.Structure):
    _fields_ = [("n_in", C.c_int32), ("n_out", C.c_int32), ("n_sliders", C.c_int32), ("kind", C.c_int32),
                ("latency", C.c_int32)]


class SliderInfo(C.Structure):
    _fields_ = [("index0", C.c_int32), ("is_choice", C.c_int32), ("def_", C.c_float), ("min", C.c_float),
                ("max", C.c_float), ("step", C.c_float), ("alias", C.c_char * 64)]


# every symbol include/za_gpu.h declares (tests check the library exports all of them)
SYMBOLS = [
    "za_gpu_plugin_count", "za_gpu_plugin_key", "za_gpu_plugin_info", "za_gpu_slider_info", "za_gpu_parse_sliders",
    "za_gpu_quantise", "za_gpu_create", "za_gpu_destroy", "za_gpu_set_stream", "za_gpu_prepare", "za_gpu_process",
    "za_gpu_process_tap64", "za_gpu_read_state", "za_gpu_sync", "za_gpu_launch_count", "za_gpu_last_kernel_ms",
    "za_gpu_set_option", "za_gpu_render", "za_gpu_last_error", "za_gpu_device_count", "za_gpu_version",
    "za_gpu_synth", "za_gpu_synth_tables", "za_gpu_checksum", "za_gpu_checksum_noalloc", "za_gpu_host_alloc", "za_gpu_host_free",
    "za_gpu_state_size", "za_gpu_state_save", "za_gpu_state_load",
    "za_gpu_create_os", "za_chain_create", "za_chain_destroy", "za_chain_info", "za_chain_stage", "za_chain_set_chunk_blocks", "za_chain_prepare",
    "za_chain_process", "za_chain_sync",
    "za_batch_create", "za_batch_destroy", "za_batch_shards", "za_batch_shard", "za_batch_prepare", "za_batch_process",
]

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ZaError(f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` "
                          "(there is no CPU fallback)")
        L = C.CDLL(str(LIB_PATH))
        L.za_gpu_plugin_key.restype = C.c_char_p
        L.za_gpu_plugin_key.argtypes = [C.c_int]
        L.za_gpu_last_error.restype = C.c_char_p
        L.za_gpu_version.restype = C.c_char_p
        L.za_gpu_plugin_info.argtypes = [C.c_char_p, C.POINTER(PluginInfo)]
        L.za_gpu_slider_info.argtypes = [C.c_char_p, C.c_int, C.POINTER(SliderInfo)]
        L.za_gpu_parse_sliders.argtypes = [C.c_char_p, C.POINTER(SliderInfo), C.c_int]
        L.za_gpu_quantise.argtypes = [C.POINTER(SliderInfo), C.c_float, C.POINTER(C.c_double)]
        L.za_gpu_create.argtypes = [C.c_char_p, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.za_gpu_create_os.argtypes = [C.c_char_p, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.za_gpu_destroy.argtypes = [C.c_void_p]
        L.za_gpu_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.za_gpu_prepare.argtypes = [C.c_void_p, C.c_void_p]
        L.za_gpu_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int]
        L.za_gpu_process_tap64.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                           C.c_int64, C.c_int]
        L.za_gpu_read_state.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        L.za_gpu_sync.argtypes = [C.c_void_p]
        L.za_gpu_launch_count.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.za_gpu_last_kernel_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.za_gpu_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.za_gpu_render.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64]
        L.za_gpu_state_size.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
        L.za_gpu_state_save.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.za_gpu_state_load.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.za_batch_create.argtypes = [C.c_char_p, C.c_int, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.za_batch_destroy.argtypes = [C.c_void_p]
        L.za_batch_shards.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.za_batch_shard.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.za_batch_prepare.argtypes = [C.c_void_p, C.c_void_p]
        L.za_batch_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.za_chain_create.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.za_chain_destroy.argtypes = [C.c_void_p]
        L.za_chain_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.za_chain_stage.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.za_chain_set_chunk_blocks.argtypes = [C.c_void_p, C.c_int]
        L.za_chain_prepare.argtypes = [C.c_void_p, C.c_void_p]
        L.za_chain_process.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int]
        L.za_chain_sync.argtypes = [C.c_void_p]
        L.za_gpu_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
        L.za_gpu_host_free.argtypes = [C.c_void_p]
        L.za_gpu_synth.argtypes = [C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p,
                                   C.c_int, C.c_void_p]
        L.za_gpu_synth_tables.argtypes = [C.c_void_p, C.c_void_p]
        L.za_gpu_checksum.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.za_gpu_checksum_noalloc.argtypes = L.za_gpu_checksum.argtypes
        _lib = L
    return _lib


def _check(rc: int, what: str):
    if rc != 0:
        raise ZaError(f"{what} failed ({rc}): {lib().za_gpu_last_error().decode()}")


def plugin_keys():
    L = lib()
    return [L.za_gpu_plugin_key(i).decode() for i in range(L.za_gpu_plugin_count())]


def plugin_info(key: str) -> PluginInfo:
    info = PluginInfo()
    _check(lib().za_gpu_plugin_info(key.encode(), C.byref(info)), "za_gpu_plugin_info")
    return info


def slider_layout(key: str):
    n = plugin_info(key).n_sliders
    out = []
    for j in range(n):
        s = SliderInfo()
        _check(lib().za_gpu_slider_info(key.encode(), j, C.byref(s)), "za_gpu_slider_info")
        out.append(s)
    return out


def parse_sliders(text: str):
    arr = (SliderInfo * 64)()
    n = lib().za_gpu_parse_sliders(text.encode(), arr, 64)
    if n < 0:
        raise ZaError("za_gpu_parse_sliders")
    return [arr[i] for i in range(n)]


def quantise(slider: SliderInfo, value: float) -> float:
    out = C.c_double()
    _check(lib().za_gpu_quantise(C.byref(slider), C.c_float(value), C.byref(out)), "za_gpu_quantise")
    return out.value


class HostBuffer:
    """Page-locked float32 host array [shape] from za_gpu_host_alloc (exact size; torch's pinned allocator rounds up to a
    power of two).  `.array` is the numpy view; free() or garbage collection releases it."""

    def __init__(self, shape):
        self.shape = tuple(int(v) for v in shape)
        n = int(np.prod(self.shape))
        self._p = C.c_void_p()
        _check(lib().za_gpu_host_alloc(n * 4, C.byref(self._p)), "za_gpu_host_alloc")
        self.array = np.ctypeslib.as_array(C.cast(self._p, C.POINTER(C.c_float)), shape=(n,)).reshape(self.shape)

    def free(self):
        if self._p:
            self.array = None
            lib().za_gpu_host_free(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def synth_into(out, kind: int, seed: int, first_stream: int = 0, t0: int = 0, cuda_stream=None):
    """Fill the CUDA tensor out[n_streams][n_ch][n_frames] with the library's device generator (za_gpu_synth)."""
    ns, nch, nf = out.shape
    assert out.is_cuda and out.is_contiguous()
    _check(lib().za_gpu_synth(kind, seed, first_stream, ns, nch, nf, nf, t0, out.data_ptr(), out.device.index or 0, cuda_stream),
           "za_gpu_synth")
    return out


def checksum(x, cuda_stream=None):
    """Per-row (stream, channel) order-independent checksum (uint64) and energy (float64) of a CUDA tensor [ns][nch][nf]."""
    ns, nch, nf = x.shape
    assert x.is_cuda and x.is_contiguous()
    s, e = np.zeros(ns * nch, np.uint64), np.zeros(ns * nch, np.float64)
    _check(lib().za_gpu_checksum(x.data_ptr(), ns * nch, nf, nf, s.ctypes.data, e.ctypes.data, x.device.index or 0, cuda_stream),
           "za_gpu_checksum")
    return s.reshape(ns, nch), e.reshape(ns, nch)


CHECKSUM_NOALLOC_MAX_ROWS = 16384


def checksum_noalloc(x, cuda_stream=None):
    """checksum() through za_gpu_checksum_noalloc: accumulators in the library's static device array (no device allocation);
    at most CHECKSUM_NOALLOC_MAX_ROWS rows per call, calls on one device must not overlap."""
    ns, nch, nf = x.shape
    assert x.is_cuda and x.is_contiguous()
    s, e = np.zeros(ns * nch, np.uint64), np.zeros(ns * nch, np.float64)
    _check(lib().za_gpu_checksum_noalloc(x.data_ptr(), ns * nch, nf, nf, s.ctypes.data, e.ctypes.data, x.device.index or 0, cuda_stream),
           "za_gpu_checksum_noalloc")
    return s.reshape(ns, nch), e.reshape(ns, nch)


def _ptr(a):
    """numpy array, torch tensor or None -> (address, keepalive)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()      # torch tensor (device or pinned host)


class BatchProcessor:
    """n_streams instances of one plugin on one B200 (one za_gpu handle)."""

    def __init__(self, plugin_key: str, n_streams: int, sample_rate: float = 48000.0, block_len: int = 512,
                 precision: str = "f64", device: int = 0, oversampling: int = 1):
        self.key = plugin_key
        self.info = plugin_info(plugin_key)
        self.n_streams = n_streams
        self.block_len = block_len
        self.precision = ZA_F64 if precision in ("f64", ZA_F64) else ZA_F32
        self._h = C.c_void_p()
        _check(lib().za_gpu_create_os(plugin_key.encode(), n_streams, sample_rate, block_len, self.precision, device, oversampling,
                                      C.byref(self._h)), "za_gpu_create")

    # prepareToPlay
    def prepare(self, param_block0=None):
        p = None if param_block0 is None else np.ascontiguousarray(param_block0, np.float32)
        if p is not None:
            assert p.shape == (self.n_streams, self.info.n_sliders), p.shape
        _check(lib().za_gpu_prepare(self._h, _ptr(p)), "za_gpu_prepare")
        return self

    # processBlock over many host blocks
    def process(self, x, out=None, params=None, tap64=None, n_frames=None, frame_stride=None):
        """x/out: numpy (host) or torch CUDA tensors [n_streams, n_ch, frames] float32.
        params: None | [n_streams, n_sliders] | [n_blocks, n_streams, n_sliders] float32 (host)."""
        is_np = isinstance(x, np.ndarray)
        dev = (not is_np) and x.is_cuda
        if n_frames is None:
            n_frames = x.shape[-1]
        if frame_stride is None:
            frame_stride = x.shape[-1]
        if out is None:
            if is_np:
                out = np.empty((self.n_streams, self.info.n_out, x.shape[-1]), np.float32)
            else:
                import torch
                out = torch.empty((self.n_streams, self.info.n_out, x.shape[-1]), dtype=torch.float32, device=x.device,
                                  pin_memory=(not dev and x.is_pinned()))
        p, npb = None, 0
        if params is not None:
            p = np.ascontiguousarray(params, np.float32)
            if p.ndim == 2:
                p = p[None]
            assert p.shape[1:] == (self.n_streams, self.info.n_sliders), p.shape
            npb = p.shape[0]
        mem = ZA_MEM_DEVICE if dev else ZA_MEM_HOST
        if tap64 is None:
            _check(lib().za_gpu_process(self._h, _ptr(p), npb, _ptr(x), _ptr(out), n_frames, frame_stride, mem),
                   "za_gpu_process")
        else:
            _check(lib().za_gpu_process_tap64(self._h, _ptr(p), npb, _ptr(x), _ptr(out), _ptr(tap64), n_frames,
                                              frame_stride, mem), "za_gpu_process_tap64")
        return out

    def read_state(self, name: str) -> np.ndarray:
        out = np.empty(self.n_streams, np.float64)
        _check(lib().za_gpu_read_state(self._h, name.encode(), out.ctypes.data), "za_gpu_read_state")
        return out

    def save_state(self) -> np.ndarray:
        """Checkpoint: the whole per-stream DSP state of the handle as one host blob (za_gpu_state_save)."""
        n = C.c_size_t()
        _check(lib().za_gpu_state_size(self._h, C.byref(n)), "za_gpu_state_size")
        blob = np.zeros(n.value, np.uint8)
        _check(lib().za_gpu_state_save(self._h, blob.ctypes.data, n.value), "za_gpu_state_save")
        return blob

    def load_state(self, blob: np.ndarray):
        blob = np.ascontiguousarray(blob, np.uint8)
        _check(lib().za_gpu_state_load(self._h, blob.ctypes.data, blob.size), "za_gpu_state_load")
        return self

    def set_stream(self, cuda_stream_ptr):
        _check(lib().za_gpu_set_stream(self._h, cuda_stream_ptr), "za_gpu_set_stream")

    def set_option(self, name: str, value: int):
        _check(lib().za_gpu_set_option(self._h, name.encode(), value), "za_gpu_set_option")

    def sync(self):
        _check(lib().za_gpu_sync(self._h), "za_gpu_sync")

    def launch_count(self) -> int:
        n = C.c_int64()
        _check(lib().za_gpu_launch_count(self._h, C.byref(n)), "za_gpu_launch_count")
        return n.value

    def last_kernel_ms(self):
        ms, k = C.c_double(), C.c_int()
        _check(lib().za_gpu_last_kernel_ms(self._h, C.byref(ms), C.byref(k)), "za_gpu_last_kernel_ms")
        return ms.value, k.value

    def close(self):
        if self._h:
            lib().za_gpu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Chain:
    """A serial plugin chain rendered on one B200 (za_chain_*): stage k's output feeds stage k + 1 on the device."""

    def __init__(self, plugin_keys, n_streams: int, sample_rate: float = 48000.0, block_len: int = 512, precision: str = "f64", device: int = 0):
        self.keys = list(plugin_keys)
        self.n_streams, self.block_len = n_streams, block_len
        arr = (C.c_char_p * len(self.keys))(*[k.encode() for k in self.keys])
        self._h = C.c_void_p()
        _check(lib().za_chain_create(arr, len(self.keys), n_streams, sample_rate, block_len, ZA_F64 if precision in ("f64", ZA_F64) else ZA_F32,
                                     device, C.byref(self._h)), "za_chain_create")
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        _check(lib().za_chain_info(self._h, C.byref(a), C.byref(b), C.byref(c)), "za_chain_info")
        self.n_in, self.n_out = a.value, b.value
        self.infos = [plugin_info(k) for k in self.keys]

    def set_chunk_blocks(self, blocks: int):
        _check(lib().za_chain_set_chunk_blocks(self._h, blocks), "za_chain_set_chunk_blocks")

    def prepare(self, params0=None):
        if params0 is None:
            _check(lib().za_chain_prepare(self._h, None), "za_chain_prepare")
            return self
        keep = [None if p is None else np.ascontiguousarray(p, np.float32) for p in params0]
        ptrs = (C.c_void_p * len(keep))(*[None if p is None else p.ctypes.data for p in keep])
        _check(lib().za_chain_prepare(self._h, ptrs), "za_chain_prepare")
        return self

    def process(self, x, params=None, out=None):
        """x [n_streams][chain n_in][frames] numpy (host) or torch CUDA tensor; params: per stage None or [n_blocks][n_streams][n_sliders]."""
        is_np = isinstance(x, np.ndarray)
        nf = x.shape[-1]
        if out is None:
            if is_np:
                out = np.empty((self.n_streams, self.n_out, nf), np.float32)
            else:
                import torch
                out = torch.empty((self.n_streams, self.n_out, nf), dtype=torch.float32, device=x.device)
        n = len(self.keys)
        keep = [None] * n if params is None else [None if p is None else np.ascontiguousarray(p, np.float32) for p in params]
        keep = [p[None] if p is not None and p.ndim == 2 else p for p in keep]
        ptrs = (C.c_void_p * n)(*[None if p is None else p.ctypes.data for p in keep])
        cnt = (C.c_int * n)(*[0 if p is None else p.shape[0] for p in keep])
        mem = ZA_MEM_HOST if is_np else ZA_MEM_DEVICE
        _check(lib().za_chain_process(self._h, ptrs, cnt, _ptr(x), _ptr(out), nf, nf, mem), "za_chain_process")
        return out

    def sync(self):
        _check(lib().za_chain_sync(self._h), "za_chain_sync")

    def close(self):
        if self._h:
            lib().za_chain_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiGpuBatch:
    """One batch sharded over several GPUs of the box (za_batch_*): host buffers in, host buffers out, one host thread per device."""

    def __init__(self, plugin_key: str, n_streams: int, devices, sample_rate: float = 48000.0, block_len: int = 512, precision: str = "f64",
                 oversampling: int = 1):
        self.info = plugin_info(plugin_key)
        self.n_streams = n_streams
        devs = (C.c_int * len(devices))(*devices)
        self._h = C.c_void_p()
        _check(lib().za_batch_create(plugin_key.encode(), n_streams, sample_rate, block_len, ZA_F64 if precision in ("f64", ZA_F64) else ZA_F32,
                                     devs, len(devices), oversampling, C.byref(self._h)), "za_batch_create")

    def prepare(self, param_block0=None):
        p = None if param_block0 is None else np.ascontiguousarray(param_block0, np.float32)
        _check(lib().za_batch_prepare(self._h, _ptr(p)), "za_batch_prepare")
        return self

    def process(self, x, params=None, out=None):
        x = np.ascontiguousarray(x, np.float32)
        nf = x.shape[-1]
        if out is None:
            out = np.empty((self.n_streams, self.info.n_out, nf), np.float32)
        p, npb = None, 0
        if params is not None:
            p = np.ascontiguousarray(params, np.float32)
            p = p[None] if p.ndim == 2 else p
            npb = p.shape[0]
        _check(lib().za_batch_process(self._h, _ptr(p), npb, _ptr(x), _ptr(out), nf, nf), "za_batch_process")
        return out

    def close(self):
        if self._h:
            lib().za_batch_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
