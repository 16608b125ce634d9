"""ctypes binding of libpose_b200.so - the C ABI declared in include/pose_b200.h.

This is plumbing for tests/ and bench.py (numpy arrays in, numpy arrays out).  It never imports
the CPU oracle and there is no fallback: if the CUDA library is missing or no device is present,
calls raise PoseError.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpose_b200.so")
HOST_LIB_PATH = os.path.join(_HERE, "libpose_b200_host.so")

PE_OK, PE_ERR_INVALID, PE_ERR_CUDA, PE_ERR_PARSE, PE_ERR_IO, PE_ERR_UNSUPPORTED, PE_ERR_NCCL = range(7)
ACT_NAMES = {0: "none", 1: "ReLU", 2: "ELU", 3: "PReLU"}


class PoseError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pose_b200 error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile the CUDA library in-tree with the committed Makefile (nvcc, sm_100a)."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-s", "-j8"] + (["-B"] if force else []))
    return LIB_PATH


_lib = None
fp, ip, u32p, vp = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.c_void_p

# name -> (restype, argtypes); every symbol include/pose_b200.h declares
SIGNATURES = {
    "pe_version": (C.c_int, []),
    "pe_last_error": (C.c_char_p, []),
    "pe_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "pe_launch_count": (C.c_int64, []),
    "pe_set_option": (C.c_int, [C.c_char_p, C.c_int]),
    "pe_get_option": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "pe_host_alloc": (C.c_int, [C.POINTER(vp), C.c_size_t]),
    "pe_host_free": (C.c_int, [vp]),
    "pe_dev_alloc": (C.c_int, [C.c_int, C.POINTER(vp), C.c_size_t]),
    "pe_dev_free": (C.c_int, [C.c_int, vp]),
    "pe_memcpy_h2d": (C.c_int, [C.c_int, vp, vp, C.c_size_t]),
    "pe_memcpy_d2h": (C.c_int, [C.c_int, vp, vp, C.c_size_t]),
    "pe_dev_memset": (C.c_int, [C.c_int, vp, C.c_int, C.c_size_t]),
    "pe_timer_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
    "pe_timer_start_net": (C.c_int, [vp, vp]),
    "pe_timer_stop_net": (C.c_int, [vp, vp]),
    "pe_timer_start_db": (C.c_int, [vp, vp]),
    "pe_timer_stop_db": (C.c_int, [vp, vp]),
    "pe_timer_elapsed_ms": (C.c_int, [vp, fp]),
    "pe_timer_destroy": (None, [vp]),
    "pe_net_profile": (C.c_int, [vp, C.c_int]),
    "pe_net_profile_read": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "pe_db_profile": (C.c_int, [vp, C.c_int]),
    "pe_db_profile_read": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "pe_db_share_stream": (C.c_int, [vp, vp]),
    "pe_net_create_from_prototxt": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
    "pe_net_create_from_string": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
    "pe_net_create_named": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
    "pe_net_destroy": (None, [vp]),
    "pe_net_info": (C.c_int, [vp] + [C.POINTER(C.c_int)] * 5),
    "pe_net_selected_channels": (C.c_int, [vp, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int)]),
    "pe_net_describe": (C.c_int, [vp, C.c_char_p, C.c_size_t]),
    "pe_net_num_blobs": (C.c_int, [vp, C.POINTER(C.c_int)]),
    "pe_net_blob_info": (C.c_int, [vp, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64)]),
    "pe_net_set_blob": (C.c_int, [vp, C.c_int, fp, C.c_int64]),
    "pe_net_get_blob": (C.c_int, [vp, C.c_int, fp, C.c_int64]),
    "pe_net_fill": (C.c_int, [vp, C.c_uint64, C.c_int]),
    "pe_net_save_weights": (C.c_int, [vp, C.c_char_p]),
    "pe_net_load_weights": (C.c_int, [vp, C.c_char_p]),
    "pe_compute_descriptors": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp]),
    "pe_compute_descriptors_dev": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp]),
    "pe_net_sync": (C.c_int, [vp]),
    "pe_synth_patches_dev": (C.c_int, [vp, C.c_uint64, C.c_int64, C.c_int64, C.c_int, C.c_int, vp]),
    "pe_synth_labels": (C.c_int, [C.c_uint64, C.c_int64, C.c_int64, C.c_int, fp]),
    "pe_binarize": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp]),
    "pe_binarize_dev": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp]),
    "pe_db_create": (C.c_int, [vp, C.c_int64, C.c_int, C.c_int64, vp, C.c_int64, C.c_int, C.c_int, C.POINTER(vp)]),
    "pe_db_create_dev": (C.c_int, [vp, C.c_int64, C.c_int, C.c_int64, vp, C.c_int64, C.c_int, C.c_int, C.POINTER(vp)]),
    "pe_db_create_hamming": (C.c_int, [vp, C.c_int64, C.c_int, C.c_int64, vp, C.c_int64, C.c_int, C.c_int, C.POINTER(vp)]),
    "pe_db_destroy": (None, [vp]),
    "pe_db_info": (C.c_int, [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pe_db_sync": (C.c_int, [vp]),
    "pe_knn": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_db_knn_flagged": (C.c_int, [vp, C.POINTER(C.c_int64)]),
    "pe_knn_dev": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_knn_hamming": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_knn_hamming_dev": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_histogram": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, vp, C.c_int, vp]),
    "pe_histogram_dev": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, vp, C.c_int, vp]),
    "pe_histogram_dev_noalloc": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, vp, C.c_int, vp]),
    "pe_confusion": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp]),
    "pe_knn_bootstrap": (C.c_int, [vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_knn_accuracy": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, fp, fp]),
    "pe_comm_unique_id": (C.c_int, [C.c_char_p]),
    "pe_comm_create": (C.c_int, [C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(vp)]),
    "pe_comm_destroy": (None, [vp]),
    "pe_comm_info": (C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pe_net_fp8_state": (C.c_int, [vp, C.POINTER(C.c_int)]),
    "pe_comm_p2p": (C.c_int, [vp, C.POINTER(C.c_int)]),
    "pe_comm_allgather_f32": (C.c_int, [vp, vp, C.c_int64, vp]),
    "pe_comm_allreduce_sum_f64": (C.c_int, [vp, vp, C.c_int64]),
    "pe_comm_allgather_f32_dev": (C.c_int, [vp, vp, vp, C.c_int64, vp]),
    "pe_knn_sharded": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_match_patches": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int, C.c_int, vp, vp, vp]),
    "pe_knn_sharded_dev": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_knn_hamming_sharded": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_knn_hamming_sharded_dev": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, vp, vp]),
    "pe_match_patches_dev": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int, C.c_int, vp, vp, vp]),
    "pe_histogram_sharded_dev": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, vp, C.c_int, vp]),
    "pe_confusion_sharded_dev": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp]),
    "pe_depth2normals": (C.c_int, [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, vp]),
    "pe_depth2normals_dev": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, vp]),
    "pe_sample_patches": (C.c_int, [C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int64, C.c_float, C.c_float,
                                    C.c_float, vp, vp]),
    "pe_sample_patches_dev": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int64, C.c_float,
                                        C.c_float, C.c_float, vp, vp]),
    "pe_samples_write": (C.c_int, [C.c_char_p, vp, vp, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int]),
    "pe_samples_info": (C.c_int, [C.c_char_p, C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.c_int)]),
    "pe_samples_read": (C.c_int, [C.c_char_p, C.c_int64, C.c_int64, vp, vp]),
    "pe_miner_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.POINTER(vp)]),
    "pe_miner_destroy": (None, [vp]),
    "pe_miner_pools_dev": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp)]),
    "pe_miner_sync": (C.c_int, [vp]),
    "pe_miner_share_stream": (C.c_int, [vp, vp]),
    "pe_miner_set_hard_negatives": (C.c_int, [vp, vp, vp]),
    "pe_miner_max_sim": (C.c_int, [vp, vp]),
    "pe_miner_compute_knn": (C.c_int, [vp, vp, vp, vp]),
    "pe_miner_set_knn": (C.c_int, [vp, vp, vp]),
    "pe_miner_build_batch": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, vp, vp]),
    "pe_miner_build_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, vp, vp]),
    "pe_miner_flags": (C.c_int, [vp, C.c_int, vp]),
    "pe_sphere_views": (C.c_int, [C.c_int, vp, vp, C.c_int, vp, C.c_int64, C.POINTER(C.c_int64)]),
    "pe_render_views": (C.c_int, [C.c_int, vp, vp, C.c_int, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, vp, vp]),
    "pe_render_views_tex": (C.c_int, [C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float,
                                      C.c_float, vp, vp]),
    "pe_render_views_tex_dev": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int,
                                          C.c_float, C.c_float, vp, vp]),
    "pe_template_samples": (C.c_int, [C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, C.c_int64, C.c_int, vp, C.c_int, C.c_int,
                                      C.c_float, C.c_float, C.c_float, vp, vp, vp]),
    "pe_template_samples_dev": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, C.c_int64, C.c_int, vp, C.c_int,
                                          C.c_int, C.c_float, C.c_float, C.c_float, vp, vp, vp]),
    "pe_pose_labels": (C.c_int, [vp, C.c_int64, C.c_int, vp]),
    "pe_scene_samples": (C.c_int, [C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp, vp, vp]),
    "pe_mesh_load_ply": (C.c_int, [C.c_char_p, C.POINTER(vp)]),
    "pe_mesh_info": (C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), fp]),
    "pe_mesh_data": (C.c_int, [vp, vp, vp, vp, vp]),
    "pe_mesh_destroy": (None, [vp]),
    "pe_render_views_dev": (C.c_int, [vp, C.c_int, vp, vp, C.c_int, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                      vp, vp]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PoseError(PE_ERR_CUDA, f"{LIB_PATH} is not built (run __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


_host_lib = None
HOST_SIGNATURES = {
    "pe_host_last_error": (C.c_char_p, []),
    "pe_host_compute_descriptors": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp]),
    "pe_host_compute_histogram": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, vp]),
    "pe_host_compute_confusion": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, C.c_char_p, C.c_char_p, C.c_int, vp]),
    "pe_host_compute_knn": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
    "pe_host_save_logs": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_char_p, C.c_int, C.c_float]),
    "pe_host_build_batch": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, vp, C.c_int, vp, vp, C.c_int, C.c_int,
                                      C.c_uint64, C.c_int, vp, vp, vp, vp]),
    "pe_host_create_synth_samples": (C.c_int, [C.c_char_p, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_char_p, vp, C.c_int,
                                               C.c_int, vp, vp, C.c_int64, C.POINTER(C.c_int64)]),
    "pe_host_extract_real_samples": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, C.c_char_p, C.c_char_p, vp, vp,
                                               C.c_int64, C.POINTER(C.c_int64)]),
    "pe_host_create_views": (C.c_int, [C.c_char_p, vp, C.c_int, vp, vp, C.c_int, vp, vp, vp, C.c_int64, C.POINTER(C.c_int64)]),
}


def host_lib():
    """libpose_b200_host.so: flat-array doors into the C++ mirror of networkEvaluator / networkSolver
    (include/networkevaluator_b200.hpp), for tests that drive it the way a C++ caller of the reference would"""
    global _host_lib
    if _host_lib is None:
        lib()
        if not os.path.exists(HOST_LIB_PATH):
            raise PoseError(PE_ERR_CUDA, f"{HOST_LIB_PATH} is not built (run __graft_entry__.build())")
        L = C.CDLL(HOST_LIB_PATH)
        for name, (res, args) in HOST_SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _host_lib = L
    return _host_lib


def host_check(code: int):
    if code != 0:
        raise PoseError(code, (host_lib().pe_host_last_error() or b"").decode())


def check(code: int):
    if code != PE_OK:
        raise PoseError(code, lib().pe_last_error().decode(errors="replace"))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(vp)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def device_count() -> int:
    n = C.c_int(0)
    code = lib().pe_device_count(C.byref(n))
    return n.value if code == PE_OK else 0


def launch_count() -> int:
    return int(lib().pe_launch_count())


def set_option(name: str, value: int):
    """process-wide switches: "knn_tensor", "forward_tensor" (1 = tcgen05 path, 0 = fp32 SIMT path)"""
    check(lib().pe_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    v = C.c_int(0)
    check(lib().pe_get_option(name.encode(), C.byref(v)))
    return v.value


class DeviceArray:
    """raw device allocation with a numpy-like shape/dtype tag"""

    def __init__(self, shape, dtype=np.float32, device=0):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        self.device = device
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = vp()
        check(lib().pe_dev_alloc(device, C.byref(p), max(self.nbytes, 16)))
        self.ptr = p.value

    @classmethod
    def from_host(cls, a, device=0):
        a = np.ascontiguousarray(a)
        d = cls(a.shape, a.dtype, device)
        if a.nbytes:
            check(lib().pe_memcpy_h2d(device, d.ptr, a.ctypes.data_as(vp), a.nbytes))
        return d

    def to_host(self):
        out = np.empty(self.shape, self.dtype)
        if self.nbytes:
            check(lib().pe_memcpy_d2h(self.device, out.ctypes.data_as(vp), self.ptr, self.nbytes))
        return out

    def offset(self, n_elems):
        return self.ptr + int(n_elems) * self.dtype.itemsize

    def free(self):
        if self.ptr:
            lib().pe_dev_free(self.device, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def pinned_empty(shape, dtype=np.float32):
    """numpy array over cudaMallocHost memory (kept alive by the returned array's base)"""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = vp()
    check(lib().pe_host_alloc(C.byref(p), max(n, 16)))
    buf = (C.c_char * max(n, 16)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[id(buf)] = (buf, p.value)
    return arr


_PINNED: dict = {}


class Timer:
    def __init__(self, device=0):
        p = vp()
        check(lib().pe_timer_create(device, C.byref(p)))
        self.h = p.value

    def start(self, obj):
        check((lib().pe_timer_start_net if isinstance(obj, Net) else lib().pe_timer_start_db)(self.h, obj.h))

    def stop(self, obj):
        check((lib().pe_timer_stop_net if isinstance(obj, Net) else lib().pe_timer_stop_db)(self.h, obj.h))

    def elapsed_ms(self) -> float:
        ms = C.c_float()
        check(lib().pe_timer_elapsed_ms(self.h, C.byref(ms)))
        return ms.value

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().pe_timer_destroy(self.h)
                self.h = None
        except Exception:       # interpreter shutdown
            pass


class Net:
    """replaces caffe::Net<float>(prototxt, caffe::TEST) (reference src/networksolver.cpp:286)"""

    def __init__(self, text: str | None = None, path: str | None = None, named: str | None = None, device: int = 0):
        h = vp()
        if named is not None:
            check(lib().pe_net_create_named(named.encode(), device, C.byref(h)))
        elif path is not None:
            check(lib().pe_net_create_from_prototxt(path.encode(), device, C.byref(h)))
        else:
            check(lib().pe_net_create_from_string(text.encode(), device, C.byref(h)))
        self.h = h.value
        self.device = device
        v = [C.c_int() for _ in range(5)]
        check(lib().pe_net_info(self.h, *[C.byref(x) for x in v]))
        self.batch, self.channels, self.height, self.width, self.desc_dim = [x.value for x in v]

    def close(self):
        if getattr(self, "h", None):
            lib().pe_net_destroy(self.h)
            self.h = None

    __del__ = close

    @property
    def sel(self):
        buf = (C.c_int * 16)()
        n = C.c_int()
        check(lib().pe_net_selected_channels(self.h, buf, 16, C.byref(n)))
        return [buf[i] for i in range(n.value)]

    def describe(self) -> str:
        buf = C.create_string_buffer(8192)
        check(lib().pe_net_describe(self.h, buf, len(buf)))
        return buf.value.decode()

    def blob_info(self):
        n = C.c_int()
        check(lib().pe_net_num_blobs(self.h, C.byref(n)))
        out = []
        for b in range(n.value):
            name = C.create_string_buffer(128)
            cnt = C.c_int64()
            check(lib().pe_net_blob_info(self.h, b, name, len(name), C.byref(cnt)))
            out.append((name.value.decode(), cnt.value))
        return out

    def set_blobs(self, blobs):
        for b, arr in enumerate(blobs):
            a = _f32(arr).reshape(-1)
            check(lib().pe_net_set_blob(self.h, b, a.ctypes.data_as(fp), a.size))

    def get_blobs(self):
        out = []
        for b, (_, cnt) in enumerate(self.blob_info()):
            a = np.empty(cnt, np.float32)
            check(lib().pe_net_get_blob(self.h, b, a.ctypes.data_as(fp), cnt))
            out.append(a)
        return out

    def fill(self, seed: int, mode: int = 1):
        check(lib().pe_net_fill(self.h, seed, mode))

    def save_weights(self, path):
        check(lib().pe_net_save_weights(self.h, str(path).encode()))

    def load_weights(self, path):
        check(lib().pe_net_load_weights(self.h, str(path).encode()))

    def compute_descriptors(self, patches_hwc: np.ndarray) -> np.ndarray:
        """networkEvaluator::computeDescriptors: n x H x W x C (HWC) host patches -> n x D"""
        x = _f32(patches_hwc)
        n = x.shape[0]
        c = x.shape[-1] if x.ndim == 4 else self.channels
        out = np.empty((n, self.desc_dim), np.float32)
        check(lib().pe_compute_descriptors(self.h, _ptr(x), n, c, _ptr(out)))
        return out

    def compute_descriptors_dev(self, d_patches, n, d_out, patch_offset=0, out_offset=0):
        check(lib().pe_compute_descriptors_dev(self.h, d_patches.offset(patch_offset * 4096 * self.channels) if isinstance(d_patches, DeviceArray) else d_patches,
                                               n, self.channels,
                                               d_out.offset(out_offset * self.desc_dim) if isinstance(d_out, DeviceArray) else d_out))

    def sync(self):
        check(lib().pe_net_sync(self.h))

    def fp8_state(self) -> int:
        """conv1 cross terms of the current weights: 0 not decided yet, 1 e4m3, 2 fp16 (pe_net_fp8_state)"""
        v = C.c_int(0)
        check(lib().pe_net_fp8_state(self.h, C.byref(v)))
        return v.value

    def profile(self, enable: bool):
        check(lib().pe_net_profile(self.h, int(enable)))

    def profile_read(self):
        """(ms per stage [conv1, conv2, fc], bracketed forward calls) since the last read"""
        ms = (C.c_double * 3)()
        calls = C.c_int64()
        check(lib().pe_net_profile_read(self.h, ms, C.byref(calls)))
        return list(ms), calls.value

    def synth_patches_dev(self, seed, first, n, poses_per_object, d_out, out_offset=0):
        ptr = d_out.offset(out_offset * 4096 * self.channels) if isinstance(d_out, DeviceArray) else d_out
        check(lib().pe_synth_patches_dev(self.h, seed, first, n, self.channels, poses_per_object, ptr))

    def binarize(self, desc: np.ndarray) -> np.ndarray:
        d = _f32(desc)
        n, dim = d.shape
        bits = np.empty((n, (dim + 31) // 32), np.uint32)
        check(lib().pe_binarize(self.h, _ptr(d), n, dim, _ptr(bits)))
        return bits

    def binarize_dev(self, d_desc, n, dim, d_bits):
        check(lib().pe_binarize_dev(self.h, d_desc.ptr, n, dim, d_bits.ptr))


def synth_labels(seed, first, n, poses_per_object) -> np.ndarray:
    out = np.empty((n, 8), np.float32)
    check(lib().pe_synth_labels(seed, first, n, poses_per_object, out.ctypes.data_as(fp)))
    return out


class TemplateDB:
    """replaces DescriptorMatcher::create("BruteForce") + add(DBfeats) (src/networkevaluator.cpp:193-196)"""

    def __init__(self, desc=None, labels_all=None, poses_per_object=0, row_offset=0, t_total=None, device=0,
                 codes=None, d_desc=None, t_local=None, dim=None):
        h = vp()
        lab = None if labels_all is None else _f32(labels_all)
        self.device = device
        if codes is not None:
            if isinstance(codes, DeviceArray):
                tl, words, cptr = codes.shape[0], codes.shape[1], codes.ptr
            else:
                codes = np.ascontiguousarray(codes, np.uint32)
                tl, words, cptr = codes.shape[0], codes.shape[1], _ptr(codes)
            tt = tl if t_total is None else t_total
            check(lib().pe_db_create_hamming(cptr, tl, words, row_offset, _ptr(lab), tt, poses_per_object, device, C.byref(h)))
            self.dim = words
        elif d_desc is not None:
            tt = t_local if t_total is None else t_total
            ptr = d_desc.ptr if isinstance(d_desc, DeviceArray) else d_desc
            check(lib().pe_db_create_dev(ptr, t_local, dim, row_offset, _ptr(lab), tt, poses_per_object, device, C.byref(h)))
            self.dim = dim
        else:
            d = _f32(desc)
            tl, dm = d.shape
            tt = tl if t_total is None else t_total
            check(lib().pe_db_create(_ptr(d), tl, dm, row_offset, _ptr(lab), tt, poses_per_object, device, C.byref(h)))
            self.dim = dm
        self.h = h.value

    def close(self):
        if getattr(self, "h", None):
            lib().pe_db_destroy(self.h)
            self.h = None

    __del__ = close

    def sync(self):
        check(lib().pe_db_sync(self.h))

    def share_stream(self, net):
        check(lib().pe_db_share_stream(self.h, net.h))

    def profile(self, enable: bool):
        check(lib().pe_db_profile(self.h, int(enable)))

    def profile_read(self):
        """(ms per stage [scan, select], bracketed kNN calls) since the last read"""
        ms = (C.c_double * 2)()
        calls = C.c_int64()
        check(lib().pe_db_profile_read(self.h, ms, C.byref(calls)))
        return list(ms), calls.value

    def knn(self, queries, k):
        """matcher->knnMatch(queries, matches, k): (idx nq x k, dist nq x k)"""
        q = _f32(queries)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int32)
        dist = np.empty((nq, k), np.float32)
        check(lib().pe_knn(self.h, _ptr(q), nq, k, _ptr(idx), _ptr(dist)))
        return idx, dist

    def knn_flagged(self) -> int:
        """queries of the last L2 kNN call that the certificate could not prove (re-scanned exactly)"""
        n = C.c_int64(0)
        check(lib().pe_db_knn_flagged(self.h, C.byref(n)))
        return n.value

    def knn_dev(self, d_q, nq, k, d_idx, d_dist):
        check(lib().pe_knn_dev(self.h, d_q.ptr if isinstance(d_q, DeviceArray) else d_q, nq, k, d_idx.ptr, d_dist.ptr))

    def knn_hamming(self, qcodes, k):
        q = np.ascontiguousarray(qcodes, np.uint32)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int32)
        dist = np.empty((nq, k), np.int32)
        check(lib().pe_knn_hamming(self.h, _ptr(q), nq, k, _ptr(idx), _ptr(dist)))
        return idx, dist

    def knn_hamming_dev(self, d_q, nq, k, d_idx, d_dist):
        check(lib().pe_knn_hamming_dev(self.h, d_q.ptr, nq, k, d_idx.ptr, d_dist.ptr))

    def knn_sharded(self, comm, queries, k):
        q = _f32(queries)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int32)
        dist = np.empty((nq, k), np.float32)
        check(lib().pe_knn_sharded(self.h, comm.h if comm else None, _ptr(q), nq, k, _ptr(idx), _ptr(dist)))
        return idx, dist

    def match_patches(self, net, comm, patches, k, want_desc=False):
        """fused computeKNN test pass (pe_match_patches): host patches (n, 64, 64, C) -> (idx, dist[, desc])"""
        x = _f32(patches)
        n, c = x.shape[0], x.shape[-1]
        idx = np.empty((n, k), np.int32)
        dist = np.empty((n, k), np.float32)
        desc = np.empty((n, self.dim), np.float32) if want_desc else None
        check(lib().pe_match_patches(net.h, self.h, comm.h if comm else None, _ptr(x), n, c, k, _ptr(idx), _ptr(dist),
                                     _ptr(desc) if want_desc else None))
        return (idx, dist, desc) if want_desc else (idx, dist)

    def match_patches_dev(self, net, comm, d_patches, n, k, d_idx, d_dist, d_desc=None, patch_offset=0):
        """pe_match_patches_dev: device patches -> neighbours of this rank's patches; asynchronous, forward on the net's
        stream, exchange + scan + merge on the DB's stream (sync() completes it)"""
        src = d_patches.offset(patch_offset * 4096 * net.channels) if isinstance(d_patches, DeviceArray) else d_patches
        check(lib().pe_match_patches_dev(net.h, self.h, comm.h if comm else None, src, n, net.channels, k, d_idx.ptr, d_dist.ptr,
                                         d_desc.ptr if d_desc is not None else None))

    def knn_hamming_sharded_dev(self, comm, d_q, nq, k, d_idx, d_dist):
        check(lib().pe_knn_hamming_sharded_dev(self.h, comm.h if comm else None, d_q.ptr, nq, k, d_idx.ptr, d_dist.ptr))

    def histogram_sharded_dev(self, comm, d_idx, d_qlabels, nq_local, k, rot_inv, bins):
        """histogram of the query set split over the ranks (every rank passes its own nq_local queries)"""
        rot, bn = _i32(rot_inv), _f32(bins)
        out = np.empty(bn.size + 2, np.float32)
        check(lib().pe_histogram_sharded_dev(self.h, comm.h if comm else None, d_idx.ptr, d_qlabels.ptr, nq_local, k, _ptr(rot),
                                             rot.size, _ptr(bn), bn.size, _ptr(out)))
        return out

    def confusion_sharded_dev(self, comm, d_idx, d_qlabels, nq_local, k, class_to_local, n_local, nr_test_poses):
        c2l = _i32(class_to_local)
        out = np.empty((n_local, n_local), np.float32)
        check(lib().pe_confusion_sharded_dev(self.h, comm.h if comm else None, d_idx.ptr, d_qlabels.ptr, nq_local, k, _ptr(c2l),
                                             c2l.size, n_local, nr_test_poses, _ptr(out)))
        return out

    def knn_sharded_dev(self, comm, d_q, nq, k, d_idx, d_dist):
        check(lib().pe_knn_sharded_dev(self.h, comm.h if comm else None, d_q.ptr if isinstance(d_q, DeviceArray) else d_q,
                                       nq, k, d_idx.ptr, d_dist.ptr))

    def knn_hamming_sharded(self, comm, qcodes, k):
        q = np.ascontiguousarray(qcodes, np.uint32)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int32)
        dist = np.empty((nq, k), np.int32)
        check(lib().pe_knn_hamming_sharded(self.h, comm.h if comm else None, _ptr(q), nq, k, _ptr(idx), _ptr(dist)))
        return idx, dist

    def histogram(self, knn_idx, q_labels, rot_inv, bins):
        """computeHistogram body: n_bins normalised bins + mean + median"""
        idx, ql = _i32(knn_idx), _f32(q_labels)
        rot, bn = _i32(rot_inv), _f32(bins)
        out = np.empty(bn.size + 2, np.float32)
        check(lib().pe_histogram(self.h, _ptr(idx), _ptr(ql), idx.shape[0], idx.shape[1], _ptr(rot), rot.size, _ptr(bn),
                                 bn.size, _ptr(out)))
        return out

    def histogram_dev(self, d_idx, d_qlabels, nq, k, rot_inv, bins):
        rot, bn = _i32(rot_inv), _f32(bins)
        out = np.empty(bn.size + 2, np.float32)
        check(lib().pe_histogram_dev(self.h, d_idx.ptr, d_qlabels.ptr, nq, k, _ptr(rot), rot.size, _ptr(bn), bn.size, _ptr(out)))
        return out

    def histogram_dev_noalloc(self, d_idx, d_qlabels, nq, k, rot_inv, bins):
        """pe_histogram_dev_noalloc: the same reduction on a fixed file-scope device scratch (no cudaMalloc); nq <= 65,536"""
        rot, bn = _i32(rot_inv), _f32(bins)
        out = np.empty(bn.size + 2, np.float32)
        check(lib().pe_histogram_dev_noalloc(self.h, d_idx.ptr, d_qlabels.ptr, nq, k, _ptr(rot), rot.size, _ptr(bn), bn.size, _ptr(out)))
        return out

    def confusion(self, knn_idx, q_labels, class_to_local, n_local, nr_test_poses):
        idx, ql, c2l = _i32(knn_idx), _f32(q_labels), _i32(class_to_local)
        out = np.empty((n_local, n_local), np.float32)
        check(lib().pe_confusion(self.h, _ptr(idx), _ptr(ql), idx.shape[0], idx.shape[1], _ptr(c2l), c2l.size, n_local,
                                 nr_test_poses, _ptr(out)))
        return out

    def knn_bootstrap(self, knn3_idx, nr_query_poses):
        idx = _i32(knn3_idx)
        nq = idx.shape[0]
        out = np.empty((nq, 4), np.int32)
        ln = np.empty(nq, np.int32)
        check(lib().pe_knn_bootstrap(self.h, _ptr(idx), nq, nr_query_poses, _ptr(out), _ptr(ln)))
        return out, ln

    def knn_accuracy(self, max_sim_tmpl0, bootstrap, nr_query_poses):
        m, b = _i32(max_sim_tmpl0), _i32(bootstrap)
        intra, inter = C.c_float(), C.c_float()
        check(lib().pe_knn_accuracy(self.h, _ptr(m), _ptr(b), m.size, nr_query_poses, C.byref(intra), C.byref(inter)))
        return intra.value, inter.value


# ----------------------------------------------------------------------- triplet miner ---
class Miner:
    """networkSolver's batch builder (reference src/networksolver.cpp:52-186, 456-505) over device-resident sample sets:
    tmpl / train patches [n_objects * poses, 64, 64, C] (or None: pools allocated, filled in place through pools()),
    labels [n_objects * poses, 8], rot_inv per used object."""

    def __init__(self, tmpl_patches, tmpl_labels, train_patches, train_labels, n_objects, rot_inv, inplane=True, channels=None,
                 device=0):
        tl, ql = _f32(tmpl_labels).reshape(-1, 8), _f32(train_labels).reshape(-1, 8)
        tp = _f32(tmpl_patches) if tmpl_patches is not None else None
        qp = _f32(train_patches) if train_patches is not None else None
        self.channels = int(channels if channels is not None else tp.shape[-1])
        self.n_objects, self.n_tmpl, self.n_train = n_objects, tl.shape[0] // n_objects, ql.shape[0] // n_objects
        ri = _i32(rot_inv)
        assert ri.size == n_objects
        h = vp()
        check(lib().pe_miner_create(device, self.channels, n_objects, self.n_tmpl, self.n_train, _ptr(tp) if tp is not None else None,
                                    _ptr(tl), _ptr(qp) if qp is not None else None, _ptr(ql), _ptr(ri), int(bool(inplane)), C.byref(h)))
        self.h = h.value
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            lib().pe_miner_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def pools(self):
        a, b = vp(), vp()
        check(lib().pe_miner_pools_dev(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def sync(self):
        check(lib().pe_miner_sync(self.h))

    def share_stream(self, net):
        check(lib().pe_miner_share_stream(self.h, net.h))

    def set_hard_negatives(self, patches, offsets):
        if offsets is None:
            check(lib().pe_miner_set_hard_negatives(self.h, None, None))
            return
        o = np.ascontiguousarray(offsets, np.int64)
        p = _f32(patches)
        check(lib().pe_miner_set_hard_negatives(self.h, _ptr(p), _ptr(o)))

    def max_sim(self):
        out = np.empty((self.n_objects * self.n_train, 2), np.int32)
        check(lib().pe_miner_max_sim(self.h, _ptr(out)))
        return out

    def compute_knn(self, net):
        nq = self.n_objects * self.n_train
        t, ln = np.empty((nq, 4), np.int32), np.empty(nq, np.int32)
        check(lib().pe_miner_compute_knn(self.h, net.h, _ptr(t), _ptr(ln)))
        return t, ln

    def set_knn(self, table, length):
        if table is None:
            check(lib().pe_miner_set_knn(self.h, None, None))
            return
        t, ln = _i32(table), _i32(length)
        check(lib().pe_miner_set_knn(self.h, _ptr(t), _ptr(ln)))

    def build_batch(self, batch_size, it, seed, triplet_size=5, random_background=0, want_batch=True):
        n = batch_size // triplet_size if triplet_size > 0 else 0
        batch = np.empty((n * 5, self.channels, 64, 64), np.float32) if want_batch else None
        idx = np.empty((n, 5), np.int32)
        check(lib().pe_miner_build_batch(self.h, batch_size, triplet_size, it, seed, random_background,
                                         _ptr(batch) if want_batch else None, _ptr(idx)))
        return batch, idx

    def build_batch_dev(self, batch_size, it, seed, d_batch, d_idx=None, triplet_size=5, random_background=0):
        check(lib().pe_miner_build_batch_dev(self.h, batch_size, triplet_size, it, seed, random_background,
                                             d_batch.ptr if isinstance(d_batch, DeviceArray) else d_batch,
                                             d_idx.ptr if isinstance(d_idx, DeviceArray) else d_idx))

    def flags(self, n_triplets):
        out = np.empty(n_triplets, np.int32)
        check(lib().pe_miner_flags(self.h, n_triplets, _ptr(out)))
        return out


# ------------------------------------------------------------------ patch extraction ---
def depth2normals(depth, fx, fy, ox=0.0, oy=0.0, device=0):
    """depth2normals (src/helper.cpp:53-103) on [n_frames, rows, cols] (or [rows, cols]) f32 depth -> normals [..., 3]"""
    d = _f32(depth)
    d3 = d.reshape((-1,) + d.shape[-2:])
    out = np.empty(d3.shape + (3,), np.float32)
    check(lib().pe_depth2normals(device, _ptr(d3), d3.shape[0], d3.shape[1], d3.shape[2], fx, fy, ox, oy, _ptr(out)))
    return out.reshape(d.shape + (3,))


def sample_patches(color, depth, centers_xy, z, fx, fy, m=0.2, normals=None, frame_of=None, device=0):
    """samplePatchWithScale (src/datasetmanager.cpp:299-346) for n patches from host frames:
    color [F, rows, cols, 3] u8, depth [F, rows, cols] f32 -> (patches [n, 64, 64, 7] f32, status [n])"""
    col = np.ascontiguousarray(color, np.uint8)
    dep = _f32(depth)
    ctr, zz = _i32(centers_xy).reshape(-1, 2), _f32(z).reshape(-1)
    nor = None if normals is None else _f32(normals)
    fr = None if frame_of is None else _i32(frame_of)
    n = ctr.shape[0]
    out = np.empty((n, 64, 64, 7), np.float32)
    status = np.empty(n, np.int32)
    check(lib().pe_sample_patches(device, _ptr(col), _ptr(dep), _ptr(nor), dep.shape[0], dep.shape[1], dep.shape[2], _ptr(fr),
                                  _ptr(ctr), _ptr(zz), n, fx, fy, m, _ptr(out), _ptr(status)))
    return out, status


def sample_patches_dev(net, d_color, d_depth, d_normals, n_frames, rows, cols, d_frame_of, d_centers, d_z, n, fx, fy, m, d_out,
                       d_status=None, device=0):
    """device-resident variant on the net's stream: the output feeds Net.compute_descriptors_dev directly"""
    p = lambda a: None if a is None else a.ptr
    check(lib().pe_sample_patches_dev(net.h if net is not None else None, device, p(d_color), p(d_depth), p(d_normals), n_frames, rows,
                                      cols, p(d_frame_of), p(d_centers), p(d_z), n, fx, fy, m, p(d_out), p(d_status)))


# ------------------------------------------------------------------ template renderer ---
def sphere_views(depth, scale, rotation, skip_lower_hemi=True):
    """SphereRenderer::createViews' poses (src/sphere.cpp:69-94): scale / rotation = (first, step, last) -> [n, 3, 4]"""
    sc, ro = _f32(scale), _f32(rotation)
    n = C.c_int64()
    check(lib().pe_sphere_views(depth, _ptr(sc), _ptr(ro), int(skip_lower_hemi), None, 0, C.byref(n)))
    out = np.empty((n.value, 3, 4), np.float32)
    check(lib().pe_sphere_views(depth, _ptr(sc), _ptr(ro), int(skip_lower_hemi), _ptr(out), n.value, C.byref(n)))
    return out


def render_views(vertices, colors, faces, cam4, poses, rows=480, cols=640, z_near=0.001, z_far=4.0, device=0):
    """Painter::paint for a batch of poses of one mesh -> (BGR u8 [n, rows, cols, 3], depth f32 [n, rows, cols])"""
    v, c, f = _f32(vertices), _f32(colors), _i32(faces)
    k, p = _f32(cam4), _f32(poses).reshape(-1, 12)
    n = p.shape[0]
    col = np.empty((n, rows, cols, 3), np.uint8)
    dep = np.empty((n, rows, cols), np.float32)
    check(lib().pe_render_views(device, _ptr(v), _ptr(c), v.shape[0], _ptr(f), f.shape[0], _ptr(k), _ptr(p), n, rows, cols, z_near, z_far,
                                _ptr(col), _ptr(dep)))
    return col, dep


def render_views_tex(vertices, colors, uv, faces, tex, cam4, poses, rows=480, cols=640, z_near=0.001, z_far=4.0, device=0):
    """textured mesh: uv [nv, 2], tex [th, tw, 3] BGR u8 (pe_render_views_tex)"""
    v, c, t, f = _f32(vertices), _f32(colors), _f32(uv), _i32(faces)
    w = np.ascontiguousarray(tex, np.uint8)
    k, p = _f32(cam4), _f32(poses).reshape(-1, 12)
    n = p.shape[0]
    col = np.empty((n, rows, cols, 3), np.uint8)
    dep = np.empty((n, rows, cols), np.float32)
    check(lib().pe_render_views_tex(device, _ptr(v), _ptr(c), _ptr(t), v.shape[0], _ptr(f), f.shape[0], _ptr(w), w.shape[0], w.shape[1],
                                    _ptr(k), _ptr(p), n, rows, cols, z_near, z_far, _ptr(col), _ptr(dep)))
    return col, dep


def pose_labels(poses, index):
    p = _f32(poses).reshape(-1, 12)
    out = np.empty((p.shape[0], 8), np.float32)
    check(lib().pe_pose_labels(_ptr(p), p.shape[0], int(index), _ptr(out)))
    return out


def template_samples(vertices, colors, faces, cam4, poses, index, uv=None, tex=None, z_override=None, rows=480, cols=640, z_near=0.001,
                     z_far=4.0, m=0.2, device=0):
    """datasetManager::createSynthSamples*: (patches [n, 64, 64, 7], labels [n, 8], status [n]) of n poses of one mesh"""
    v, c, f = _f32(vertices), _f32(colors), _i32(faces)
    k, p = _f32(cam4), _f32(poses).reshape(-1, 12)
    t = _f32(uv) if uv is not None else None
    w = np.ascontiguousarray(tex, np.uint8) if tex is not None else None
    z = _f32(z_override) if z_override is not None else None
    n = p.shape[0]
    patches = np.empty((n, 64, 64, 7), np.float32)
    labels = np.empty((n, 8), np.float32)
    status = np.empty(n, np.int32)
    check(lib().pe_template_samples(device, _ptr(v), _ptr(c), _ptr(t), v.shape[0], _ptr(f), f.shape[0], _ptr(w), w.shape[0] if w is not None else 0,
                                    w.shape[1] if w is not None else 0, _ptr(k), _ptr(p), n, int(index), _ptr(z), rows, cols, z_near, z_far, m,
                                    _ptr(patches), _ptr(labels), _ptr(status)))
    return patches, labels, status


def scene_samples(color, depth, cam4, poses, index, centroid=None, z_from_depth=False, m=0.2, device=0):
    """datasetManager::extractRealSamplesPaul / Wadim: frames + ground-truth poses -> (patches, labels, status)"""
    col, dep = np.ascontiguousarray(color, np.uint8), _f32(depth)
    k, p = _f32(cam4), _f32(poses).reshape(-1, 12)
    c = _f32(centroid) if centroid is not None else None
    n, rows, cols = dep.shape
    patches = np.empty((n, 64, 64, 7), np.float32)
    labels = np.empty((n, 8), np.float32)
    status = np.empty(n, np.int32)
    check(lib().pe_scene_samples(device, _ptr(col), _ptr(dep), n, rows, cols, _ptr(k), _ptr(p), _ptr(c), int(z_from_depth), int(index), m,
                                 _ptr(patches), _ptr(labels), _ptr(status)))
    return patches, labels, status


def load_ply(path):
    """Model::loadModel: dict(vertices [nv, 3], colors [nv, 3] in [0, 1], uv [nv, 2] or None, faces [nf, 3], centroid [3])"""
    h = vp()
    check(lib().pe_mesh_load_ply(str(path).encode(), C.byref(h)))
    try:
        nv, nf, hc, hu = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        cen = np.zeros(3, np.float32)
        check(lib().pe_mesh_info(h, C.byref(nv), C.byref(nf), C.byref(hc), C.byref(hu), cen.ctypes.data_as(fp)))
        v, c = np.empty((nv.value, 3), np.float32), np.empty((nv.value, 3), np.float32)
        uv = np.empty((nv.value, 2), np.float32) if hu.value else None
        f = np.empty((nf.value, 3), np.int32)
        check(lib().pe_mesh_data(h, _ptr(v), _ptr(c), _ptr(uv) if uv is not None else None, _ptr(f)))
        return dict(vertices=v, colors=c, uv=uv, faces=f, centroid=cen, has_colors=bool(hc.value))
    finally:
        lib().pe_mesh_destroy(h)


def render_views_dev(net, d_vertices, d_colors, n_vertices, d_faces, n_faces, cam4, d_poses, n_views, rows, cols, d_color, d_depth,
                     z_near=0.001, z_far=4.0, device=0):
    k = _f32(cam4)
    check(lib().pe_render_views_dev(net.h if net is not None else None, device, d_vertices.ptr, d_colors.ptr, n_vertices, d_faces.ptr,
                                    n_faces, _ptr(k), d_poses.ptr, n_views, rows, cols, z_near, z_far, d_color.ptr, d_depth.ptr))


# ------------------------------------------------------------------ sample-set files ---
def samples_write(path, data_hwc, labels):
    """hdf5Handler::write (src/hdf5handler.cpp:71-132) into a flat PESS0001 file: data [n, H, W, C], labels [n, L]"""
    d, l = _f32(data_hwc), _f32(labels)
    n, hh, ww, cc = d.shape
    check(lib().pe_samples_write(os.fsencode(path), _ptr(d), _ptr(l), n, cc, hh, ww, l.shape[1] if l.ndim == 2 else 0))


def samples_info(path):
    n, c, hh, ww, ld = C.c_int64(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
    check(lib().pe_samples_info(os.fsencode(path), C.byref(n), C.byref(c), C.byref(hh), C.byref(ww), C.byref(ld)))
    return {"n": n.value, "channels": c.value, "height": hh.value, "width": ww.value, "label_dim": ld.value}


def samples_read(path, first=0, count=None):
    """hdf5Handler::read (src/hdf5handler.cpp:11-69) -> (data [count, H, W, C], labels [count, L])"""
    i = samples_info(path)
    count = i["n"] - first if count is None else count
    d = np.empty((count, i["height"], i["width"], i["channels"]), np.float32)
    l = np.empty((count, i["label_dim"]), np.float32)
    check(lib().pe_samples_read(os.fsencode(path), first, count, _ptr(d), _ptr(l)))
    return d, l


class Comm:
    """NCCL communicator for the row-sharded DB (one process per GPU)"""

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib().pe_comm_unique_id(buf))
        return buf.raw

    def __init__(self, rank, world, uid: bytes, device):
        h = vp()
        check(lib().pe_comm_create(rank, world, uid, device, C.byref(h)))
        self.h = h.value
        self.rank, self.world = rank, world

    def p2p_active(self) -> bool:
        """True when the all-gathers run as one peer-memory kernel (raises if one of them timed out)"""
        v = C.c_int(0)
        check(lib().pe_comm_p2p(self.h, C.byref(v)))
        return bool(v.value)

    def allgather_f32(self, a):
        a = _f32(a).reshape(-1)
        out = np.empty((self.world, a.size), np.float32)
        check(lib().pe_comm_allgather_f32(self.h, _ptr(a), a.size, _ptr(out)))
        return out

    def allgather_f32_dev(self, net, d_send, n, d_recv):
        check(lib().pe_comm_allgather_f32_dev(self.h, net.h, d_send.ptr, n, d_recv.ptr))

    def allreduce_sum_f64(self, a):
        a = np.ascontiguousarray(a, np.float64).copy()
        check(lib().pe_comm_allreduce_sum_f64(self.h, _ptr(a), a.size))
        return a

    def close(self):
        if getattr(self, "h", None):
            lib().pe_comm_destroy(self.h)
            self.h = None
