"""ctypes binding of oracle/_ref/libnrc_ref.so -- the UNMODIFIED reference cache (nrc::Network + vendored
tiny-cuda-nn) compiled for sm_100 by oracle/ref/Makefile.ref, wrapped by oracle/ref/ref_wrap.cu.
TEST INFRASTRUCTURE ONLY: used to generate tests/golden/ fixtures on a GPU box, by GPU parity tests, and by
`bench.py --ref-gpu` to time the reference on the same box.  Never imported by the product."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_ref", "libnrc_ref.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, i32, f32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_float
        L.nrc_ref_last_error.restype = ctypes.c_char_p
        L.nrc_ref_create.argtypes = [i32, vp, ctypes.POINTER(vp)]
        L.nrc_ref_destroy.argtypes = [vp]
        L.nrc_ref_infer.argtypes = [vp, vp, vp, u32, vp]
        L.nrc_ref_train.argtypes = [vp, vp, vp, vp, ctypes.POINTER(f32)]
        L.nrc_ref_set_lr.argtypes = [vp, f32]
        L.nrc_ref_n_params.argtypes = [vp]
        L.nrc_ref_get_params.argtypes = [vp, i32, vp]
        L.nrc_ref_set_master.argtypes = [vp, vp]
        if hasattr(L, "nrc_refh_accumulate"):    # the reference's frame helper kernels (ref_helpers_wrap.cu)
            L.nrc_refh_last_error.restype = ctypes.c_char_p
            L.nrc_refh_accumulate.argtypes = [vp, vp, vp, i32, i32, i32, i32]
            L.nrc_refh_propagate.argtypes = [vp, vp, vp, vp, i32, i32]
            L.nrc_refh_permute.argtypes = [vp, vp, vp, vp, vp, i32]
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise RuntimeError("reference: " + lib().nrc_ref_last_error().decode())


class RefNetwork:
    """nrc::Network of the reference, driven on the current CUDA device.  Buffers are raw device pointers."""

    def __init__(self, encoding, stream=0):
        h = ctypes.c_void_p()
        _ck(lib().nrc_ref_create(int(encoding), stream, ctypes.byref(h)))
        self.h, self.stream = h, stream
        self.n = lib().nrc_ref_n_params(self.h)

    def destroy(self):
        if self.h:
            _ck(lib().nrc_ref_destroy(self.h))
            self.h = None

    def infer(self, in_d, out_d, n):
        """NOTE: the reference rounds n up to a multiple of 256 and touches those rows (NRCNetwork.cu:72)."""
        _ck(lib().nrc_ref_infer(self.h, in_d, out_d, n, self.stream))

    def train(self, in_d, tgt_d, want_loss=True):
        loss = ctypes.c_float()
        _ck(lib().nrc_ref_train(self.h, in_d, tgt_d, self.stream, ctypes.byref(loss) if want_loss else None))
        return loss.value if want_loss else None

    def set_lr(self, lr):
        _ck(lib().nrc_ref_set_lr(self.h, lr))

    def get_params(self, which):
        """0 = fp32 master, 1 = fp16 training weights, 2 = fp16 inference (EMA) weights, 3 = fp16 gradients."""
        out = np.empty(self.n, dtype=np.float32 if which == 0 else np.float16)
        _ck(lib().nrc_ref_get_params(self.h, which, out.ctypes.data))
        return out

    def set_master(self, master):
        m = np.ascontiguousarray(master, dtype=np.float32)
        assert m.size == self.n
        _ck(lib().nrc_ref_set_master(self.h, m.ctypes.data))


# ---- the reference's frame helper kernels (nrc/shaders/nrc_helpers.cu:77-249), unmodified source; raw device pointers
def _hck(rc):
    if rc != 0:
        raise RuntimeError("reference helpers: " + lib().nrc_refh_last_error().decode())


def accumulate_render_radiance(radiance_d, throughput_d, output_d, width, height, mode, iteration):
    _hck(lib().nrc_refh_accumulate(radiance_d, throughput_d, output_d, width, height, mode, iteration))


def propagate_train_radiance(end_vertices_d, end_radiance_d, records_d, targets_d, tiles_x, tiles_y):
    _hck(lib().nrc_refh_propagate(end_vertices_d, end_radiance_d, records_d, targets_d, tiles_x, tiles_y))


def permute_train_data(q_src_d, q_dst_d, t_src_d, t_dst_d, permutation_d, num_records):
    _hck(lib().nrc_refh_permute(q_src_d, q_dst_d, t_src_d, t_dst_d, permutation_d, num_records))
