"""Host-side mirror of the reference operator interface for the hot path.

Same names, argument meaning and error behaviour as FFTW_Convolution in
Convolution/src/convolution_fftw.h: ``Workspace`` (:29-42), ``init_workspace`` (:44-148),
``convolve`` (:216-284), ``clear_workspace`` (:150-164).  ``Plan`` is the batched, device-resident
form the reference does not have.  Everything calls the C-ABI; nothing is computed in Python.
"""
import ctypes

import numpy as np

from . import capi
from .capi import check, lib


def find_closest_factor(n):
    """The reference planner (factorize.h:92-104 with FFTW_FACTORS), restated in the library."""
    return lib().b2c_find_closest_factor(int(n))


def smooth_size(n):
    """The new planner: cheapest 2/3/5/7-smooth period >= n for the shared-memory radix schedule."""
    return lib().b2c_smooth_size(int(n))


class Plan:
    """One (mode, h_src, w_src, h_kernel, w_kernel) problem: wraps a b2c_plan*."""

    def __init__(self, mode, h_src, w_src, h_kernel, w_kernel):
        self._h = ctypes.c_void_p()
        check(lib().b2c_plan_create(ctypes.byref(self._h), int(mode), int(h_src), int(w_src),
                                    int(h_kernel), int(w_kernel)))
        self.mode = int(mode)
        self.h_src, self.w_src, self.h_kernel, self.w_kernel = int(h_src), int(w_src), int(h_kernel), int(w_kernel)
        a, b, c, d = (ctypes.c_int() for _ in range(4))
        check(lib().b2c_plan_query(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c), ctypes.byref(d)))
        self.h_fft, self.w_fft, self.h_dst, self.w_dst = a.value, b.value, c.value, d.value

    # -- planning facts
    def info(self):
        path, oh, ow, fh, fw, sm = (ctypes.c_int() for _ in range(6))
        wb = ctypes.c_long()
        check(lib().b2c_plan_info(self._h, ctypes.byref(path), ctypes.byref(oh), ctypes.byref(ow),
                                  ctypes.byref(fh), ctypes.byref(fw), ctypes.byref(wb), ctypes.byref(sm)))
        return dict(path=path.value, off_h=oh.value, off_w=ow.value, fold_h=fh.value, fold_w=fw.value,
                    work_bytes=wb.value, work_in_smem=sm.value)

    def set_path(self, path):
        check(lib().b2c_plan_set_path(self._h, int(path)))

    # -- compute, host buffers (numpy float64, C-contiguous)
    def convolve_host(self, src, kernel):
        src = np.ascontiguousarray(src, dtype=np.float64)
        kernel = np.ascontiguousarray(kernel, dtype=np.float64)
        assert src.shape == (self.h_src, self.w_src) and kernel.shape == (self.h_kernel, self.w_kernel)
        dst = np.zeros((self.h_dst, self.w_dst), dtype=np.float64)
        check(lib().b2c_convolve_host(self._h, src.ctypes.data, kernel.ctypes.data, dst.ctypes.data))
        return dst

    def convolve_batch_host(self, src, kernel, dst=None):
        src = np.ascontiguousarray(src, dtype=np.float64)
        kernel = np.ascontiguousarray(kernel, dtype=np.float64)
        n = src.shape[0]
        assert src.shape[1:] == (self.h_src, self.w_src)
        shared = 1 if kernel.ndim == 2 else 0
        assert kernel.shape[-2:] == (self.h_kernel, self.w_kernel) and (shared or kernel.shape[0] == n)
        if dst is None:
            dst = np.zeros((n, self.h_dst, self.w_dst), dtype=np.float64)
        check(lib().b2c_convolve_batch_host(self._h, src.ctypes.data, kernel.ctypes.data, shared,
                                            dst.ctypes.data, n))
        return dst

    def convolve_batch_host_ptr(self, src_ptr, kernel_ptr, shared, dst_ptr, n):
        check(lib().b2c_convolve_batch_host(self._h, src_ptr, kernel_ptr, int(shared), dst_ptr, int(n)))

    # -- compute, device buffers (raw pointers: torch tensors' data_ptr(), stream handle or 0)
    def convolve_batch_device(self, d_src, d_kernel, shared, d_dst, n, stream=0):
        check(lib().b2c_convolve_batch_device(self._h, int(d_src), int(d_kernel), int(shared), int(d_dst),
                                              int(n), ctypes.c_void_p(int(stream))))

    def convolve_batch_device_f32(self, d_src, d_kernel, shared, d_dst, n, stream=0):
        """float32 device buffers in and out, FP64 arithmetic (b2c_convolve_batch_device_f32)."""
        check(lib().b2c_convolve_batch_device_f32(self._h, int(d_src), int(d_kernel), int(shared), int(d_dst),
                                                  int(n), ctypes.c_void_p(int(stream))))

    def convolve_batch_device_c64(self, d_src, d_kernel, shared, d_dst, n, stream=0):
        """complex128 (interleaved) device buffers, complex kernel (b2c_convolve_batch_device_c64)."""
        check(lib().b2c_convolve_batch_device_c64(self._h, int(d_src), int(d_kernel), int(shared), int(d_dst),
                                                  int(n), ctypes.c_void_p(int(stream))))

    def destroy(self):
        if self._h:
            check(lib().b2c_plan_destroy(self._h))
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def multi_convolve_batch_host(mode, src, kernel, n_devices=0):
    """Shard a host-resident batch over the GPUs of this process (b2c_multi_convolve_batch_host): one
    host thread, plan and pipeline per device, contiguous shards, no inter-GPU traffic."""
    src = np.ascontiguousarray(src, dtype=np.float64)
    kernel = np.ascontiguousarray(kernel, dtype=np.float64)
    n, hs, ws = src.shape
    shared = 1 if kernel.ndim == 2 else 0
    hk, wk = kernel.shape[-2:]
    assert shared or kernel.shape[0] == n
    p = Plan(mode, hs, ws, hk, wk)
    hd, wd = p.h_dst, p.w_dst
    p.destroy()
    dst = np.zeros((n, hd, wd), dtype=np.float64)
    L = lib()
    L.b2c_multi_convolve_batch_host.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                                    ctypes.c_void_p, ctypes.c_long, ctypes.c_int]
    check(L.b2c_multi_convolve_batch_host(int(mode), hs, ws, hk, wk, src.ctypes.data, kernel.ctypes.data, shared,
                                          dst.ctypes.data if dst.size else None, n, int(n_devices)))
    return dst


class Workspace:
    """Mirror of FFTW_Convolution::Workspace (convolution_fftw.h:29-42): the fields callers read
    are ``dst``, ``h_dst``, ``w_dst``; ``h_fftw``/``w_fftw`` hold the new planner's periods."""

    def __init__(self):
        self.plan = None
        self.h_src = self.w_src = self.h_kernel = self.w_kernel = 0
        self.h_fftw = self.w_fftw = 0
        self.mode = None
        self.dst = None
        self.h_dst = self.w_dst = 0


def init_workspace(ws, mode, h_src, w_src, h_kernel, w_kernel):
    """convolution_fftw.h:44-148.  An unknown mode prints the list of modes and carries on, like
    the reference (:122-131); the workspace is then left without a plan and convolve() is a no-op."""
    ws.h_src, ws.w_src, ws.h_kernel, ws.w_kernel, ws.mode = h_src, w_src, h_kernel, w_kernel, mode
    try:
        ws.plan = Plan(mode, h_src, w_src, h_kernel, w_kernel)
    except capi.B2CError as e:
        if e.code != 1:
            raise
        print(str(e))
        ws.plan = None
        ws.h_fftw = ws.w_fftw = ws.h_dst = ws.w_dst = 0
        ws.dst = None
        return
    ws.h_fftw, ws.w_fftw, ws.h_dst, ws.w_dst = ws.plan.h_fft, ws.plan.w_fft, ws.plan.h_dst, ws.plan.w_dst
    ws.dst = np.zeros((ws.h_dst, ws.w_dst), dtype=np.float64)


def convolve(ws, src, kernel):
    """convolution_fftw.h:216-284: result in ws.dst (h_dst x w_dst); empty plan -> no-op (:218-219)."""
    if ws.plan is None or ws.h_fftw <= 0 or ws.w_fftw <= 0:
        return
    src = np.ascontiguousarray(src, dtype=np.float64).reshape(ws.h_src, ws.w_src)
    kernel = np.ascontiguousarray(kernel, dtype=np.float64).reshape(ws.h_kernel, ws.w_kernel)
    check(lib().b2c_convolve_host(ws.plan._h, src.ctypes.data, kernel.ctypes.data, ws.dst.ctypes.data))


def clear_workspace(ws):
    """convolution_fftw.h:150-164."""
    if ws.plan is not None:
        ws.plan.destroy()
        ws.plan = None
    ws.dst = None
