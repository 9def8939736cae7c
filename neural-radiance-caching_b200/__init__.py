"""nrc_b200 -- Python binding of the B200-native neural radiance cache (C ABI in include/nrc_b200.h).

The directory name carries a hyphen (it mirrors the reference repo's name), so import it through
``__graft_entry__.load_package()`` which registers it as ``nrc_b200``.

``Network`` mirrors the reference's cache interface ``nrc::Network``
(/root/reference/nrc/inc/NRCNetwork.h:20-75): same method names, argument meaning and error behaviour
(exceptions instead of status codes, silent no-ops after ``destroy``).  Device buffers are raw device
pointers (ints) or torch CUDA tensors; torch is only used for memory and streams.

There is NO CPU path: if libnrc_b200.so is missing, or the device is not sm_100, construction raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnrc_b200.so")
# A/B measurements only (tools/ab_variants.py): another BUILD of the same CUDA library, e.g. build/variants/libnrc_b200_v0.so
if os.environ.get("NRC_B200_LIB"):
    LIB_PATH = os.path.abspath(os.environ["NRC_B200_LIB"])

FREQUENCY, HASH = 0, 1            # nrc::InputEncoding (neural_radiance_caching.h:24-27)
NUM_BATCHES = 4                   # neural_radiance_caching.h:29
BATCH_SIZE = 16384                # neural_radiance_caching.h:31
NN_INPUT_DIMS, NN_OUTPUT_DIMS = 15, 3

PARAMS_MASTER_F32, PARAMS_TRAIN_F16, PARAMS_EMA_F16, PARAMS_GRAD_F16, ADAM_M1_F32, ADAM_M2_F32, GRAD_F32 = range(7)
_F32_KINDS = (PARAMS_MASTER_F32, ADAM_M1_F32, ADAM_M2_F32, GRAD_F32)

# every symbol include/nrc_b200.h declares (tests check the library exports all of them)
ABI_SYMBOLS = [
    "nrc_b200_create", "nrc_b200_create_ex", "nrc_b200_destroy", "nrc_b200_reset", "nrc_b200_infer",
    "nrc_b200_train", "nrc_b200_set_stream", "nrc_b200_set_learning_rate", "nrc_b200_get_learning_rate",
    "nrc_b200_infer_host", "nrc_b200_train_host", "nrc_b200_last_loss_device", "nrc_b200_train_grad", "nrc_b200_train_apply",
    "nrc_b200_n_params", "nrc_b200_get_params", "nrc_b200_set_params", "nrc_b200_load_master",
    "nrc_b200_last_error", "nrc_b200_launch_count", "nrc_b200_debug_train_timeline", "nrc_b200_dp_export", "nrc_b200_dp_connect", "nrc_b200_train_dp",
    "nrc_b200_accumulate_render_radiance", "nrc_b200_propagate_train_radiance", "nrc_b200_permutation",
    "nrc_b200_permute_train_data", "nrc_b200_train_shuffled", "nrc_b200_train_batches", "nrc_b200_adjust_tile_size", "nrc_b200_update",
    "nrc_b200_debug_l2_read_bandwidth", "nrc_b200_release", "nrc_b200_set_accumulator", "nrc_b200_dp_connect_local", "nrc_b200_permute_train_data_with",
    "nrc_b200_train_grad_noalloc", "nrc_b200_noalloc_limits",
]
ACC_FP32, ACC_FP16 = 0, 1         # nrc_b200_set_accumulator


# nrc::RenderMode (neural_radiance_caching.h:13-22)
RENDER_FULL, RENDER_NO_CACHE, RENDER_CACHE_ONLY, RENDER_CACHE_FIRST_VERTEX, RENDER_DEBUG_NO_THROUGHPUT, RENDER_DEBUG_THROUGHPUT = range(6)


class Frame(ctypes.Structure):       # nrc_b200_frame
    _fields_ = [("queries_d", ctypes.c_void_p), ("results_d", ctypes.c_void_p), ("num_pixels", ctypes.c_uint32),
                ("num_tiles", ctypes.c_uint32), ("throughput_d", ctypes.c_void_p), ("output_rgba_d", ctypes.c_void_p),
                ("render_mode", ctypes.c_int), ("iteration_index", ctypes.c_uint32), ("end_vertices_d", ctypes.c_void_p),
                ("records_d", ctypes.c_void_p), ("train_queries_d", ctypes.c_void_p), ("train_targets_d", ctypes.c_void_p),
                ("num_train_records", ctypes.c_int), ("shuffle_seed", ctypes.c_uint32)]


class TrainingStat(ctypes.Structure):   # nrc::TrainingStat, NRCNetwork.h:15-18
    _fields_ = [("loss", ctypes.c_float), ("num_train_records", ctypes.c_int)]


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into libnrc_b200.so (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc")], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libnrc_b200.so failed")
    return LIB_PATH


_lib = None


def lib():
    """The loaded C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no fallback path)")
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, u64, f32, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_float, ctypes.c_int
        P = ctypes.POINTER
        L.nrc_b200_create.argtypes = [i32, vp, P(vp)]
        L.nrc_b200_create_ex.argtypes = [i32, i32, u64, vp, P(vp)]
        L.nrc_b200_destroy.argtypes = [vp]
        L.nrc_b200_reset.argtypes = [vp, i32]
        L.nrc_b200_infer.argtypes = [vp, vp, vp, u32, vp]
        L.nrc_b200_train.argtypes = [vp, vp, vp, vp, P(f32)]
        L.nrc_b200_set_stream.argtypes = [vp, vp]
        L.nrc_b200_set_learning_rate.argtypes = [vp, f32]
        L.nrc_b200_get_learning_rate.argtypes = [vp, P(f32)]
        L.nrc_b200_infer_host.argtypes = [vp, vp, vp, u32]
        L.nrc_b200_train_host.argtypes = [vp, vp, vp, P(f32)]
        L.nrc_b200_last_loss_device.argtypes = [vp, P(vp)]
        L.nrc_b200_train_grad.argtypes = [vp, vp, vp, u32, u32, P(vp), vp]
        L.nrc_b200_train_apply.argtypes = [vp, vp, P(f32)]
        if hasattr(L, "nrc_b200_train_grad_noalloc"):   # absent only from older A/B builds selected with NRC_B200_LIB
            L.nrc_b200_train_grad_noalloc.argtypes = [vp, vp, vp, u32, u32, vp, vp, vp]
            L.nrc_b200_noalloc_limits.argtypes = [P(u32), P(ctypes.c_uint64)]
        L.nrc_b200_n_params.argtypes = [vp, P(u32), P(u32)]
        L.nrc_b200_get_params.argtypes = [vp, i32, vp]
        L.nrc_b200_set_params.argtypes = [vp, i32, vp]
        L.nrc_b200_load_master.argtypes = [vp, vp]
        L.nrc_b200_last_error.restype = ctypes.c_char_p
        L.nrc_b200_launch_count.argtypes = [vp, P(u64)]
        L.nrc_b200_debug_train_timeline.argtypes = [vp, i32, vp]
        L.nrc_b200_accumulate_render_radiance.argtypes = [vp, vp, vp, u32, i32, u32, vp]
        L.nrc_b200_propagate_train_radiance.argtypes = [vp, vp, vp, vp, u32, vp]
        L.nrc_b200_permutation.argtypes = [u32, u32]
        L.nrc_b200_permutation.restype = u32
        L.nrc_b200_permute_train_data.argtypes = [vp, vp, vp, vp, i32, u32, vp]
        L.nrc_b200_train_shuffled.argtypes = [vp, vp, vp, i32, u32, u32, vp, P(f32)]
        L.nrc_b200_train_batches.argtypes = [vp, vp, vp, i32, vp, P(f32)]
        L.nrc_b200_adjust_tile_size.argtypes = [i32, i32, i32, P(i32), P(i32)]
        L.nrc_b200_adjust_tile_size.restype = None
        L.nrc_b200_update.argtypes = [vp, P(Frame), vp, P(TrainingStat)]
        L.nrc_b200_dp_export.argtypes = [vp, vp]
        L.nrc_b200_dp_connect.argtypes = [vp, i32, i32, vp]
        L.nrc_b200_train_dp.argtypes = [vp, vp, vp, u32, u32, vp, P(f32)]
        L.nrc_b200_release.argtypes = [vp]
        L.nrc_b200_debug_l2_read_bandwidth.argtypes = [u64, i32, P(f32)]
        L.nrc_b200_set_accumulator.argtypes = [vp, i32]
        L.nrc_b200_dp_connect_local.argtypes = [P(vp), i32]
        L.nrc_b200_permute_train_data_with.argtypes = [vp, vp, vp, vp, vp, i32, vp]
        _lib = L
    return _lib


def _ptr(x):
    """Raw device/host pointer of a torch tensor, numpy array, ctypes pointer or int."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    return ctypes.cast(x, ctypes.c_void_p).value


def _check(rc):
    """True when the call ran; False for a call on a destroyed network (-3: silent no-op, NRCNetwork.cu:43,66)."""
    if rc == 0:
        return True
    if rc == -3:
        return False
    msg = lib().nrc_b200_last_error().decode()
    if rc == -1:
        raise ValueError(msg)    # std::invalid_argument (NRCNetworkConfigs.h:129-131)
    raise RuntimeError(msg)      # std::runtime_error (CUDA_CHECK_THROW)


def accumulate_render_radiance(radiance_d, throughput_d, output_rgba_d, n_pixels, render_mode, iteration_index, stream=None):
    _check(lib().nrc_b200_accumulate_render_radiance(_ptr(radiance_d), _ptr(throughput_d), _ptr(output_rgba_d), int(n_pixels),
                                                     int(render_mode), int(iteration_index), _ptr(stream)))


def propagate_train_radiance(end_vertices_d, end_radiance_d, records_d, targets_d, n_tiles, stream=None):
    _check(lib().nrc_b200_propagate_train_radiance(_ptr(end_vertices_d), _ptr(end_radiance_d), _ptr(records_d),
                                                   _ptr(targets_d), int(n_tiles), _ptr(stream)))


def permute_train_data(q_src_d, t_src_d, q_dst_d, t_dst_d, num_records, seed, stream=None):
    _check(lib().nrc_b200_permute_train_data(_ptr(q_src_d), _ptr(t_src_d), _ptr(q_dst_d), _ptr(t_dst_d), int(num_records),
                                             int(seed), _ptr(stream)))


def permute_train_data_with(q_src_d, t_src_d, q_dst_d, t_dst_d, permutation_d, num_records, stream=None):
    _check(lib().nrc_b200_permute_train_data_with(_ptr(q_src_d), _ptr(t_src_d), _ptr(q_dst_d), _ptr(t_dst_d), _ptr(permutation_d),
                                                  int(num_records), _ptr(stream)))


def dp_connect_local(networks):
    """Connects the data-parallel exchange between Network objects of THIS process (one per GPU), rank = list index."""
    arr = (ctypes.c_void_p * len(networks))(*[n._h for n in networks])
    _check(lib().nrc_b200_dp_connect_local(arr, len(networks)))


def l2_read_bandwidth(nbytes=32 << 20, reps=20):
    """Measured GB/s the SMs read from an L2-resident buffer (bench.py's gather-roofline denominator)."""
    v = ctypes.c_float()
    _check(lib().nrc_b200_debug_l2_read_bandwidth(int(nbytes), int(reps), ctypes.byref(v)))
    return v.value


def permutation(i, seed):
    return lib().nrc_b200_permutation(int(i), int(seed))


def adjust_tile_size(tile_x, tile_y, num_train_records):
    x, y = ctypes.c_int(), ctypes.c_int()
    lib().nrc_b200_adjust_tile_size(int(tile_x), int(tile_y), int(num_train_records), ctypes.byref(x), ctypes.byref(y))
    return x.value, y.value


class HyperParams:               # nrc::HyperParams, NRCNetwork.h:10-13
    def __init__(self, learningRate):
        self.learningRate = learningRate


class Network:
    """Mirror of ``nrc::Network``.  ``init`` must be called before use, ``destroy`` explicitly at the end."""

    def __init__(self):
        self._h = None
        self._stream = None

    # template<bool Verbose> void init(CUstream, nrc::InputEncoding)     NRCNetwork.h:26-31
    def init(self, stream, encoding, verbose=False, log2_hashmap_size=15, seed=1337):
        if self._h is not None:              # init on a live (or destroyed) object = full re-create (NRCNetwork.cu:106-112)
            _check(lib().nrc_b200_release(self._h))
            self._h = None
        h = ctypes.c_void_p()
        _check(lib().nrc_b200_create_ex(int(encoding), int(log2_hashmap_size), int(seed), _ptr(stream), ctypes.byref(h)))
        self._h, self._stream = h, stream
        if verbose:
            print(f"nrc_b200 network: encoding={'Frequency' if encoding == 0 else 'Hash'} n_params={self.n_params()}")

    def destroy(self):                       # NRCNetwork.cu:35-39
        """Frees the device memory; the handle stays behind as a tombstone (later calls are silent no-ops, as in the
        reference) until the object is re-initialised or garbage-collected."""
        if self._h is not None:
            _check(lib().nrc_b200_destroy(self._h))

    def __del__(self):
        try:
            if self._h is not None:
                lib().nrc_b200_release(self._h)
                self._h = None
        except Exception:
            pass

    def set_accumulator(self, mode):
        """ACC_FP32 (default) / ACC_FP16 (the reference's fp16-accumulator arithmetic) for the query kernel."""
        _check(lib().nrc_b200_set_accumulator(self._h, int(mode)))

    def train(self, batchInputs_d, batchTargets_d, stream=None, want_loss=False):   # NRCNetwork.cu:41-62
        """One optimizer step on BATCH_SIZE samples.  Returns the loss (host float, forces a stream sync as the
        reference's loss_h != nullptr path does) when want_loss, else None (asynchronous)."""
        if self._h is None:
            return None
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train(self._h, _ptr(batchInputs_d), _ptr(batchTargets_d), _ptr(stream),
                                    ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    def infer(self, inputs_d, outputs_d, numInputs, stream=None):                   # NRCNetwork.cu:64-83
        if self._h is None:
            return
        _check(lib().nrc_b200_infer(self._h, _ptr(inputs_d), _ptr(outputs_d), int(numInputs), _ptr(stream)))

    def setStream(self, stream):             # NRCNetwork.cu:85-88
        if self._h is not None:
            _check(lib().nrc_b200_set_stream(self._h, _ptr(stream)))

    def setHyperParams(self, hp):            # NRCNetwork.cu:90-94
        _check(lib().nrc_b200_set_learning_rate(self._h, float(hp.learningRate)))

    def getLearningRate(self):               # NRCNetwork.cu:101-104
        v = ctypes.c_float()
        _check(lib().nrc_b200_get_learning_rate(self._h, ctypes.byref(v)))
        return v.value

    # ---- additions (no reference counterpart) ---------------------------------------------------------
    def infer_host(self, inputs_h, outputs_h, numInputs):
        _check(lib().nrc_b200_infer_host(self._h, _ptr(inputs_h), _ptr(outputs_h), int(numInputs)))

    def train_host(self, inputs_h, targets_h, want_loss=False):
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train_host(self._h, _ptr(inputs_h), _ptr(targets_h),
                                         ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    def train_grad(self, inputs_d, targets_d, n_local, n_global, stream=None):
        """Gradient of this rank's shard; returns the device pointer of the fp32 gradient (n_params floats)."""
        g = ctypes.c_void_p()
        _check(lib().nrc_b200_train_grad(self._h, _ptr(inputs_d), _ptr(targets_d), int(n_local), int(n_global),
                                         ctypes.byref(g), _ptr(stream)))
        return g.value

    def train_grad_noalloc(self, inputs_d, targets_d, n_local, n_global, grad_out_d, loss_out_d=None, stream=None):
        """train_grad with its scratch in the library's static device array: allocates nothing; the fp32 gradient (n_params
        floats) and the loss go to caller buffers.  n_local <= noalloc_limits()[0].  No counterpart in the reference."""
        _check(lib().nrc_b200_train_grad_noalloc(self._h, _ptr(inputs_d), _ptr(targets_d), int(n_local), int(n_global),
                                                 _ptr(grad_out_d), _ptr(loss_out_d), _ptr(stream)))

    def train_apply(self, stream=None, want_loss=False):
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train_apply(self._h, _ptr(stream), ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    def dp_export(self):
        """64-byte CUDA IPC handle of this rank's exchange buffer (bytes)."""
        buf = ctypes.create_string_buffer(64)
        _check(lib().nrc_b200_dp_export(self._h, buf))
        return buf.raw

    def dp_connect(self, rank, world, handles):
        """handles: bytes, world x 64, in rank order (all-gathered by the caller)."""
        assert len(handles) == 64 * world
        _check(lib().nrc_b200_dp_connect(self._h, int(rank), int(world), ctypes.c_char_p(handles)))

    def train_dp(self, inputs_d, targets_d, n_local, n_global, stream=None, want_loss=False):
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train_dp(self._h, _ptr(inputs_d), _ptr(targets_d), int(n_local), int(n_global), _ptr(stream),
                                       ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    # ---- NEXT rows: the per-frame sequence around the network (Device::render) ------------------------------
    def train_batches(self, queries_d, targets_d, n_batches, stream=None, want_loss=False):
        """n_batches consecutive train() calls (Device::nrcTrainRadiance) as one launch; returns the mean loss if asked."""
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train_batches(self._h, _ptr(queries_d), _ptr(targets_d), int(n_batches), _ptr(stream),
                                            ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    def train_shuffled(self, q_src_d, t_src_d, num_records, seed, batch_index, stream=None, want_loss=False):
        loss = ctypes.c_float()
        ran = _check(lib().nrc_b200_train_shuffled(self._h, _ptr(q_src_d), _ptr(t_src_d), int(num_records), int(seed),
                                             int(batch_index), _ptr(stream), ctypes.byref(loss) if want_loss else None))
        return loss.value if (want_loss and ran) else None

    def update(self, frame, stream=None, want_stat=False):
        """frame: nrc_b200.Frame.  Returns TrainingStat when want_stat (one synchronisation), else None."""
        stat = TrainingStat()
        _check(lib().nrc_b200_update(self._h, ctypes.byref(frame), _ptr(stream), ctypes.byref(stat) if want_stat else None))
        return stat if want_stat else None

    def last_loss_device(self):
        p = ctypes.c_void_p()
        _check(lib().nrc_b200_last_loss_device(self._h, ctypes.byref(p)))
        return p.value

    def n_params(self):
        a, b = ctypes.c_uint32(), ctypes.c_uint32()
        _check(lib().nrc_b200_n_params(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def get_params(self, kind):
        import numpy as np
        n, _ = self.n_params()
        dt = np.float32 if kind in _F32_KINDS else np.float16
        out = np.empty(n, dtype=dt)
        _check(lib().nrc_b200_get_params(self._h, kind, out.ctypes.data))
        return out

    def set_params(self, kind, arr):
        import numpy as np
        n, _ = self.n_params()
        dt = np.float32 if kind in _F32_KINDS else np.float16
        a = np.ascontiguousarray(arr, dtype=dt)
        assert a.size == n
        _check(lib().nrc_b200_set_params(self._h, kind, a.ctypes.data))

    def load_master(self, master):
        import numpy as np
        a = np.ascontiguousarray(master, dtype=np.float32)
        assert a.size == self.n_params()[0]
        _check(lib().nrc_b200_load_master(self._h, a.ctypes.data))

    def debug_train_timeline(self, enable=True, read=False):
        import numpy as np
        out = np.zeros((128, 8), dtype=np.int64) if read else None
        _check(lib().nrc_b200_debug_train_timeline(self._h, int(enable), out.ctypes.data if read else None))
        return out

    def launch_count(self):
        c = ctypes.c_uint64()
        _check(lib().nrc_b200_launch_count(self._h, ctypes.byref(c)))
        return c.value


def noalloc_limits():
    """(max samples per nrc_b200_train_grad_noalloc call, bytes of the static device scratch every GPU holds)."""
    n, b = ctypes.c_uint32(), ctypes.c_uint64()
    _check(lib().nrc_b200_noalloc_limits(ctypes.byref(n), ctypes.byref(b)))
    return n.value, b.value
