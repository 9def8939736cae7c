"""ctypes binding of the CPU ORACLE (oracle/nrc_oracle.c).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs -- never by the product."""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnrc_oracle.so")


def build():
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


class Spec(ctypes.Structure):
    _fields_ = [("encoding", ctypes.c_int), ("enc_width", ctypes.c_int), ("n_mlp_params", ctypes.c_int),
                ("n_grid_params", ctypes.c_int), ("n_params", ctypes.c_int), ("n_levels", ctypes.c_int),
                ("log2_hashmap", ctypes.c_int), ("level_offset", ctypes.c_uint32 * 17)]


class State(ctypes.Structure):
    _fields_ = [("spec", Spec), ("master", ctypes.POINTER(ctypes.c_float)), ("params", ctypes.POINTER(ctypes.c_uint16)),
                ("ema", ctypes.POINTER(ctypes.c_uint16)), ("grads", ctypes.POINTER(ctypes.c_uint16)),
                ("m1", ctypes.POINTER(ctypes.c_float)), ("m2", ctypes.POINTER(ctypes.c_float)),
                ("steps", ctypes.POINTER(ctypes.c_uint32)), ("step", ctypes.c_uint32),
                ("lr", ctypes.c_float), ("beta1", ctypes.c_float), ("beta2", ctypes.c_float), ("eps", ctypes.c_float),
                ("l2_reg", ctypes.c_float), ("ema_decay", ctypes.c_float), ("loss_scale", ctypes.c_float)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, i32, f32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_float
        L.nrc_oracle_make_spec.argtypes = [i32, i32, ctypes.POINTER(Spec)]
        L.nrc_oracle_encode.argtypes = [ctypes.POINTER(Spec), vp, vp, u32, vp]
        L.nrc_oracle_infer.argtypes = [ctypes.POINTER(Spec), vp, vp, u32, vp, i32, i32]
        L.nrc_oracle_state_create.argtypes = [i32, i32]
        L.nrc_oracle_state_create.restype = ctypes.POINTER(State)
        L.nrc_oracle_state_destroy.argtypes = [ctypes.POINTER(State)]
        L.nrc_oracle_state_init.argtypes = [ctypes.POINTER(State), ctypes.c_uint64]
        L.nrc_oracle_state_set_master.argtypes = [ctypes.POINTER(State), vp, i32]
        L.nrc_oracle_train_step.argtypes = [ctypes.POINTER(State), vp, vp, u32, i32]
        L.nrc_oracle_train_step.restype = f32
        L.nrc_oracle_grad_only.argtypes = [ctypes.POINTER(State), vp, vp, u32, u32, vp, i32]
        L.nrc_oracle_grad_only.restype = f32
        L.nrc_oracle_apply_grads.argtypes = [ctypes.POINTER(State), vp]
        L.nrc_oracle_set_grid_grad_mode.argtypes = [i32]
        L.nrc_oracle_set_train_acc_mode.argtypes = [i32]
        L.nrc_oracle_quartic_cdf.argtypes = [f32, f32]
        L.nrc_oracle_quartic_cdf.restype = f32
        L.nrc_oracle_oneblob4_aos.argtypes = [f32, ctypes.POINTER(f32 * 4)]
        L.nrc_oracle_oneblob4_soa.argtypes = [f32, ctypes.POINTER(f32 * 4)]
        L.nrc_oracle_triwave.argtypes = [f32, i32]
        L.nrc_oracle_triwave.restype = f32
        L.nrc_oracle_grid_index.argtypes = [u32] * 5
        L.nrc_oracle_grid_index.restype = u32
        L.nrc_oracle_accumulate.argtypes = [vp, vp, vp, u32, i32, u32]
        L.nrc_oracle_propagate.argtypes = [vp, vp, vp, vp, u32]
        L.nrc_oracle_permutation.argtypes = [u32, u32]
        L.nrc_oracle_permutation.restype = u32
        L.nrc_oracle_permute.argtypes = [vp, vp, vp, vp, i32, u32]
        L.nrc_oracle_adjust_tile_size.argtypes = [i32, i32, i32, ctypes.POINTER(i32), ctypes.POINTER(i32)]
        _lib = L
    return _lib


def set_grid_grad_mode(mode):
    """0 (default): Hash grid gradients summed in fp32 (what the CUDA kernels do by default); 1: emulation of the
    reference's arithmetic -- half products added with fp16 atomics (grid.h:252-255,301-319), in sample order."""
    lib().nrc_oracle_set_grid_grad_mode(int(mode))


def set_train_acc_mode(mode):
    """Accumulators of the TRAINING forward pass: 0 (default) fp32; 1 fp16 per 16-wide k block, the arithmetic of the
    reference's wmma::accumulator<__half> fragments (fully_fused_mlp.cu:68)."""
    lib().nrc_oracle_set_train_acc_mode(int(mode))


def make_spec(encoding, log2_hashmap=15):
    s = Spec()
    if lib().nrc_oracle_make_spec(encoding, log2_hashmap, ctypes.byref(s)) != 0:
        raise ValueError("Unsupported input encoding")
    return s


def ncpu():
    return os.cpu_count() or 1


def encode(spec, params_h, queries):
    """params_h: float16 array of n_params; returns float16 [n, E]."""
    q = np.ascontiguousarray(queries, dtype=np.float32)
    p = np.ascontiguousarray(params_h, dtype=np.float16)
    out = np.empty((q.shape[0], spec.enc_width), dtype=np.float16)
    lib().nrc_oracle_encode(ctypes.byref(spec), p.ctypes.data, q.ctypes.data, q.shape[0], out.ctypes.data)
    return out


def infer(spec, params_h, queries, acc_mode=0, nthreads=None):
    q = np.ascontiguousarray(queries, dtype=np.float32)
    p = np.ascontiguousarray(params_h, dtype=np.float16)
    assert p.size == spec.n_params
    out = np.empty((q.shape[0], 3), dtype=np.float32)
    lib().nrc_oracle_infer(ctypes.byref(spec), p.ctypes.data, q.ctypes.data, q.shape[0], out.ctypes.data, acc_mode,
                           nthreads or ncpu())
    return out


class Trainer:
    """Oracle of tcnn::Trainer as nrc::Network drives it (train = training_step + loss)."""

    def __init__(self, encoding, log2_hashmap=15, seed=1337):
        self.st = lib().nrc_oracle_state_create(encoding, log2_hashmap)
        if not self.st:
            raise ValueError("Unsupported input encoding")
        lib().nrc_oracle_state_init(self.st, seed)
        self.n = self.st.contents.spec.n_params

    def close(self):
        if self.st:
            lib().nrc_oracle_state_destroy(self.st)
            self.st = None

    @property
    def spec(self):
        return self.st.contents.spec

    def _arr(self, ptr, dtype):
        return np.ctypeslib.as_array(ptr, shape=(self.n,)).view(dtype)

    master = property(lambda s: s._arr(s.st.contents.master, np.float32))
    params = property(lambda s: s._arr(s.st.contents.params, np.float16))
    ema = property(lambda s: s._arr(s.st.contents.ema, np.float16))
    grads = property(lambda s: s._arr(s.st.contents.grads, np.float16))
    m1 = property(lambda s: s._arr(s.st.contents.m1, np.float32))
    m2 = property(lambda s: s._arr(s.st.contents.m2, np.float32))

    def set_lr(self, lr):
        self.st.contents.lr = lr

    def set_master(self, master, reset_optimizer=True):
        m = np.ascontiguousarray(master, dtype=np.float32)
        assert m.size == self.n
        lib().nrc_oracle_state_set_master(self.st, m.ctypes.data, int(reset_optimizer))

    def train_step(self, queries, targets, nthreads=None):
        q = np.ascontiguousarray(queries, dtype=np.float32)
        t = np.ascontiguousarray(targets, dtype=np.float32)
        return lib().nrc_oracle_train_step(self.st, q.ctypes.data, t.ctypes.data, q.shape[0], nthreads or ncpu())

    def grad_only(self, queries, targets, n_norm, grad_accum, nthreads=None):
        q = np.ascontiguousarray(queries, dtype=np.float32)
        t = np.ascontiguousarray(targets, dtype=np.float32)
        assert grad_accum.dtype == np.float32 and grad_accum.size == self.n
        return lib().nrc_oracle_grad_only(self.st, q.ctypes.data, t.ctypes.data, q.shape[0], n_norm,
                                          grad_accum.ctypes.data, nthreads or ncpu())

    def apply_grads(self, grad_accum):
        lib().nrc_oracle_apply_grads(self.st, grad_accum.ctypes.data)

    def infer(self, queries, use_ema=True, acc_mode=0, nthreads=None):
        return infer(self.spec, self.ema if use_ema else self.params, queries, acc_mode, nthreads)


# ---- NEXT rows (frame helpers) ---------------------------------------------------------------------------------
RECORD_DTYPE = np.dtype([("propTo", np.int32), ("localThroughput", np.float32, 3), ("pixelIndex", np.int32),
                         ("tileIndex", np.int32), ("propLength", np.int32)])                       # nrc::TrainingRecord
END_VERTEX_DTYPE = np.dtype([("startTrainRecord", np.int32), ("radianceMask", np.float32), ("pixelIndex", np.int32),
                             ("tileIndex", np.int32)])                                               # TrainingSuffixEndVertex


def accumulate(radiance, throughput, out4, mode, iteration_index):
    r = np.ascontiguousarray(radiance, np.float32); t = np.ascontiguousarray(throughput, np.float32)
    o = np.ascontiguousarray(out4, np.float32).copy()
    lib().nrc_oracle_accumulate(r.ctypes.data, t.ctypes.data, o.ctypes.data, r.shape[0], mode, iteration_index)
    return o


def propagate(end_vertices, end_radiance, records, targets):
    ev = np.ascontiguousarray(end_vertices); rec = np.ascontiguousarray(records)
    assert ev.dtype == END_VERTEX_DTYPE and rec.dtype == RECORD_DTYPE
    er = np.ascontiguousarray(end_radiance, np.float32); t = np.ascontiguousarray(targets, np.float32).copy()
    lib().nrc_oracle_propagate(ev.ctypes.data, er.ctypes.data, rec.ctypes.data, t.ctypes.data, ev.shape[0])
    return t


def permutation(i, seed):
    return lib().nrc_oracle_permutation(int(i), int(seed))


def permute(q_src, t_src, num_records, seed):
    q = np.ascontiguousarray(q_src, np.float32); t = np.ascontiguousarray(t_src, np.float32)
    qd = np.zeros((65536, 15), np.float32); td = np.zeros((65536, 3), np.float32)
    lib().nrc_oracle_permute(q.ctypes.data, t.ctypes.data, qd.ctypes.data, td.ctypes.data, int(num_records), int(seed))
    return qd, td


def adjust_tile_size(tx, ty, num_records):
    x, y = ctypes.c_int(), ctypes.c_int()
    lib().nrc_oracle_adjust_tile_size(tx, ty, num_records, ctypes.byref(x), ctypes.byref(y))
    return x.value, y.value
