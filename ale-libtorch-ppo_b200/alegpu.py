"""alegpu.py — host-side mirror of the reference's learner interfaces over libalegpu.so.

The reference (cemlyn007/ale-libtorch-ppo) exposes its learner path as C++ functions over
torch::Tensor (src/ai/gae.h, src/ai/buffer.h, src/ai/vision.h, src/ai/ppo/losses.h,
src/ai/ppo/train.h).  This module gives the same operations, with the same names, argument
order and error behaviour, over the C ABI in include/alegpu.h: `gae`, `Buffer`
(`add`/`get` -> `Batch`), `losses.compute`/`normalize_logits`, `Learner` (= Network +
torch::optim::Adam + train()), `Metrics`, `Hyperparameters`.  std::invalid_argument maps to
ValueError, std::runtime_error to RuntimeError, std::out_of_range to IndexError.

Device memory is owned by the library (cudaMalloc through the ABI); numpy arrays are staged
with explicit copies.  There is NO CPU fallback: if libalegpu.so is missing or no CUDA
device is present every call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libalegpu.so")
_lib = None

MATH_FP32, MATH_BF16, MATH_TF32 = 0, 1, 2
FRAME_BYTES = 210 * 160 * 3
STACK_BYTES = 4 * 84 * 84


class CudaError(RuntimeError):
    pass


def lib():
    """Load libalegpu.so (built by `make -C ale-libtorch-ppo_b200`); fail loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with __graft_entry__.build() "
                "(there is no CPU fallback for the learner path)")
        _lib = C.CDLL(SO_PATH)
        _lib.alegpu_last_error.restype = C.c_char_p
        _lib.alegpu_version.restype = C.c_char_p
        _lib.alegpu_learner_param_count.restype = C.c_int64
        _lib.alegpu_learner_param_count.argtypes = [C.c_void_p]
        _lib.alegpu_learner_comm_mode.restype = C.c_char_p
        _lib.alegpu_learner_comm_mode.argtypes = [C.c_void_p]
        _lib.alegpu_learner_stream.restype = C.c_void_p
        _lib.alegpu_learner_stream.argtypes = [C.c_void_p]
        _lib.alegpu_learner_last_launch_count.restype = C.c_int64
        _lib.alegpu_learner_last_launch_count.argtypes = [C.c_void_p]
        _lib.alegpu_launch_count.restype = C.c_int64
    return _lib


def _check(rc):
    if rc == 0:
        return
    msg = lib().alegpu_last_error().decode()
    if rc == 1:
        raise ValueError(msg)       # std::invalid_argument
    if rc == 2:
        raise RuntimeError(msg)     # std::runtime_error
    if rc == 3:
        raise IndexError(msg)       # std::out_of_range
    raise CudaError(msg)


def device_count():
    n = C.c_int(0)
    try:
        _check(lib().alegpu_device_count(C.byref(n)))
    except CudaError:
        return 0
    return n.value


def launch_count():
    return int(lib().alegpu_launch_count())


def synchronize(stream=None):
    _check(lib().alegpu_stream_synchronize(C.c_void_p(stream)))


class DeviceArray:
    """A typed view of library-owned device memory."""

    def __init__(self, shape, dtype, ptr=None, owner=None):
        self.shape = tuple(int(s) for s in (shape if hasattr(shape, "__len__") else (shape,)))
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize
        self._owned = ptr is None
        self._owner = owner
        if ptr is None:
            p = C.c_void_p()
            _check(lib().alegpu_malloc(C.byref(p), C.c_size_t(max(self.nbytes, 1))))
            ptr = p.value
        self.ptr = ptr

    @classmethod
    def from_numpy(cls, a, dtype=None, stream=None):
        a = np.ascontiguousarray(a, dtype=dtype)
        d = cls(a.shape, a.dtype)
        d.copy_from(a, stream)
        return d

    @classmethod
    def zeros(cls, shape, dtype):
        d = cls(shape, dtype)
        _check(lib().alegpu_memset(C.c_void_p(d.ptr), 0, C.c_size_t(d.nbytes), None))
        return d

    def copy_from(self, a, stream=None):
        a = np.ascontiguousarray(a, dtype=self.dtype)
        assert a.nbytes == self.nbytes, (a.shape, self.shape)
        _check(lib().alegpu_memcpy_h2d(C.c_void_p(self.ptr), a.ctypes.data_as(C.c_void_p),
                                       C.c_size_t(self.nbytes), C.c_void_p(stream)))
        synchronize(stream)

    def numpy(self, stream=None):
        out = np.empty(self.shape, self.dtype)
        if self.nbytes:
            _check(lib().alegpu_memcpy_d2h(out.ctypes.data_as(C.c_void_p), C.c_void_p(self.ptr),
                                           C.c_size_t(self.nbytes), C.c_void_p(stream)))
            synchronize(stream)
        return out

    def view(self, offset_elems, shape):
        """Sub-view starting `offset_elems` elements in (no copy)."""
        return DeviceArray(shape, self.dtype, self.ptr + offset_elems * self.dtype.itemsize, self)

    def __del__(self):
        if getattr(self, "_owned", False) and getattr(self, "ptr", None) and _lib is not None:
            _lib.alegpu_free(C.c_void_p(self.ptr))
            self.ptr = None


def _p(x):
    if x is None:
        return C.c_void_p(None)
    return C.c_void_p(x.ptr if isinstance(x, DeviceArray) else int(x))


def _dev(a, dtype):
    return a if isinstance(a, DeviceArray) else DeviceArray.from_numpy(a, dtype)


# ---- ai::gae::gae — src/ai/gae.h:4-7 ---------------------------------------------------------
def gae(advantages, rewards, values, next_values, terminals, truncations, episode_starts, gamma,
        lambda_, stream=None):
    """Writes `advantages` (DeviceArray [E,T]) in place, like the reference's out-parameter."""
    E, T = rewards.shape
    _check(lib().alegpu_gae(_p(advantages), _p(rewards), _p(values), _p(next_values),
                            _p(terminals), _p(truncations), _p(episode_starts), C.c_int64(E),
                            C.c_int64(T), C.c_float(gamma), C.c_float(lambda_),
                            C.c_void_p(stream)))


def gae_numpy(rewards, values, next_values, terminals, truncations, episode_starts, gamma, lam):
    r, v, nv = (_dev(x, np.float32) for x in (rewards, values, next_values))
    te, tr, st = (_dev(x, np.uint8) for x in (terminals, truncations, episode_starts))
    adv = DeviceArray.zeros(r.shape, np.float32)
    gae(adv, r, v, nv, te, tr, st, gamma, lam)
    return adv.numpy()


def buffer_get_numpy(rewards, values, next_values, terminals, truncations, episode_starts, gamma,
                     lam, write_index=0):
    """Buffer::get in pointer form (src/ai/buffer.cc:58-77) on numpy inputs."""
    r, v, nv = (DeviceArray.from_numpy(x, np.float32) for x in (rewards, values, next_values))
    te, tr, st = (_dev(x, np.uint8) for x in (terminals, truncations, episode_starts))
    E, T = r.shape
    adv, ret = DeviceArray.zeros((E, T), np.float32), DeviceArray.zeros((E, T), np.float32)
    mk = DeviceArray.zeros((E, T), np.uint8)
    _check(lib().alegpu_buffer_get(_p(r), _p(v), _p(nv), _p(te), _p(tr), _p(st), C.c_int64(E),
                                   C.c_int64(T), C.c_float(gamma), C.c_float(lam),
                                   C.c_int64(write_index), _p(adv), _p(ret), _p(mk), 0, None,
                                   None))
    return dict(rewards=r.numpy(), advantages=adv.numpy(), returns=ret.numpy(), masks=mk.numpy())


# ---- preprocessing (max_and_skip.cc:33-42 + vision.cc:38-84 + rollout.cc:184-196) -----------
def preprocess_step(frames_a, frames_b, starts, src_stack, dst_stack, gray_f32=None,
                    starts_stride=1, src_env_stride=STACK_BYTES, dst_env_stride=STACK_BYTES,
                    E=None, stream=None):
    E = frames_a.shape[0] if E is None else E
    _check(lib().alegpu_preprocess_step(_p(frames_a), _p(frames_b), _p(starts),
                                        C.c_int64(starts_stride), _p(src_stack),
                                        C.c_int64(src_env_stride), _p(dst_stack),
                                        C.c_int64(dst_env_stride), _p(gray_f32), C.c_int64(E),
                                        C.c_void_p(stream)))


def preprocess_rollout(frames, episode_starts, carry_in, carry_out, obs, E, T, stream=None):
    _check(lib().alegpu_preprocess_rollout(_p(frames), _p(episode_starts), _p(carry_in),
                                           _p(carry_out), _p(obs), C.c_int64(E), C.c_int64(T),
                                           C.c_void_p(stream)))


# ---- live path: per-frame stb resize (vision.cc:86-95) -> 2-frame max -> stack ------------------
class vision:
    """ai::vision — src/ai/vision.h:15-17 (the std::vector<unsigned char> overload)."""

    @staticmethod
    def resize_grayscale_image(image, width, height, new_width, new_height, return_f32=False):
        """image: [height*width] or [n,height,width] u8 (host or DeviceArray) -> same rank, resized.
        Restated stbir_resize_uint8_linear — parity unpinned, see include/alegpu.h."""
        host = not isinstance(image, DeviceArray)
        a = np.ascontiguousarray(image, dtype=np.uint8) if host else image
        size = int(np.prod(a.shape))
        if width <= 0 or height <= 0 or size % (width * height) != 0:
            raise ValueError("image.size() must be a multiple of width*height")  # vision.cc:89
        n = size // (width * height)
        flat = len(a.shape) == 1
        x = _dev(a, np.uint8)
        out = DeviceArray((n, new_height, new_width) if new_height > 0 and new_width > 0 else (0,),
                          np.uint8)
        f32 = DeviceArray((n, new_height, new_width), np.float32) if return_f32 else None
        _check(lib().alegpu_resize_grayscale(_p(x), C.c_int64(n), C.c_int(width), C.c_int(height),
                                             C.c_int(new_width), C.c_int(new_height), _p(out),
                                             _p(f32), None))
        res = out.numpy() if host else out
        if host and flat:
            res = res.reshape(-1)
        return (res, f32.numpy()) if return_f32 else res


def vision_resize_area(planes):
    """interpolate(size 84x84, mode area) over [..., 210, 160] float planes (alegpu_vision_resize_area:
    resize_grayscale_image / resize_frame_stacked_{grayscale,rgb}_images, src/ai/vision.cc:12-48)."""
    x = np.ascontiguousarray(planes, np.float32)
    lead = x.shape[:-2]
    n = int(np.prod(lead)) if lead else 1
    d, out = DeviceArray.from_numpy(x), DeviceArray((n, 84, 84), np.float32)
    _check(lib().alegpu_vision_resize_area(_p(d), C.c_int64(n), _p(out), None))
    return out.numpy().reshape(lead + (84, 84))


def vision_rgb_to_grayscale(images):
    """rgb_to_grayscale_frame_stacked_images (src/ai/vision.cc:71-84): [..., 3, 84, 84] -> [..., 84, 84]."""
    x = np.ascontiguousarray(images, np.float32)
    lead = x.shape[:-3]
    n = int(np.prod(lead)) if lead else 1
    d, out = DeviceArray.from_numpy(x), DeviceArray((n, 84, 84), np.float32)
    _check(lib().alegpu_vision_rgb_to_grayscale(_p(d), C.c_int64(n), _p(out), None))
    return out.numpy().reshape(lead + (84, 84))


def preprocess_gray_step(frames_a, frames_b, starts, src_stack, dst_stack, frame_env_stride=210 * 160,
                         starts_stride=1, src_env_stride=STACK_BYTES, dst_env_stride=STACK_BYTES,
                         E=None, stream=None):
    """resize(a), resize(b) -> max -> stack update (rollout.cc:146-161, 184-196), one kernel."""
    E = frames_a.shape[0] if E is None else E
    _check(lib().alegpu_preprocess_gray_step(_p(frames_a), _p(frames_b),
                                             C.c_int64(frame_env_stride), _p(starts),
                                             C.c_int64(starts_stride), _p(src_stack),
                                             C.c_int64(src_env_stride), _p(dst_stack),
                                             C.c_int64(dst_env_stride), C.c_int64(E),
                                             C.c_void_p(stream)))


# ---- ai::ppo::losses — src/ai/ppo/losses.h:6-37 ----------------------------------------------
class losses:
    @staticmethod
    def normalize_logits(logits):
        z = _dev(logits, np.float32)
        out = DeviceArray(z.shape, np.float32)
        _check(lib().alegpu_normalize_logits(_p(z), C.c_int64(z.shape[0]), C.c_int64(z.shape[1]),
                                             _p(out), None))
        return out.numpy()

    @staticmethod
    def compute(logits, old_log_probabilities, actions, advantages, values, returns, masks,
                clip_param, value_loss_coef, entropy_coef):
        """losses::compute on RAW logits + the autograd gradient w.r.t. (logits, values)."""
        z, ol = _dev(logits, np.float32), _dev(old_log_probabilities, np.float32)
        a = _dev(actions, np.int64)
        adv, v, ret = (_dev(x, np.float32) for x in (advantages, values, returns))
        mk = _dev(masks, np.uint8)
        n, A = z.shape
        out = {k: DeviceArray.zeros(n, np.float32) for k in
               ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dvalues")}
        out["dlogits"] = DeviceArray.zeros((n, A), np.float32)
        loss = DeviceArray.zeros(1, np.float32)
        _check(lib().alegpu_loss(_p(z), _p(ol), _p(a), _p(adv), _p(v), _p(ret), _p(mk),
                                 C.c_int64(n), C.c_int64(A), C.c_float(clip_param),
                                 C.c_float(value_loss_coef), C.c_float(entropy_coef), _p(loss),
                                 _p(out["clipped_losses"]), _p(out["value_losses"]),
                                 _p(out["entropies"]), _p(out["total_losses"]), _p(out["ratio"]),
                                 _p(out["dlogits"]), _p(out["dvalues"]), None))
        res = {k: d.numpy() for k, d in out.items()}
        res["loss"] = float(loss.numpy()[0])
        return res

    LOSS_STATIC_MAX_ROWS = 1 << 20  # ALEGPU_LOSS_STATIC_MAX_ROWS

    @staticmethod
    def compute_static(logits, old_log_probabilities, actions, advantages, values, returns, masks,
                       clip_param, value_loss_coef, entropy_coef):
        """alegpu_loss_static: losses.compute with the library's static device scratch (no allocation
        inside the call; n <= LOSS_STATIC_MAX_ROWS).  No counterpart in the reference."""
        z, ol = _dev(logits, np.float32), _dev(old_log_probabilities, np.float32)
        a = _dev(actions, np.int64)
        adv, v, ret = (_dev(x, np.float32) for x in (advantages, values, returns))
        mk = _dev(masks, np.uint8)
        n, A = z.shape
        out = {k: DeviceArray.zeros(n, np.float32) for k in
               ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dvalues")}
        out["dlogits"] = DeviceArray.zeros((n, A), np.float32)
        loss = DeviceArray.zeros(1, np.float32)
        _check(lib().alegpu_loss_static(_p(z), _p(ol), _p(a), _p(adv), _p(v), _p(ret), _p(mk),
                                        C.c_int64(n), C.c_int64(A), C.c_float(clip_param),
                                        C.c_float(value_loss_coef), C.c_float(entropy_coef),
                                        _p(loss), _p(out["clipped_losses"]),
                                        _p(out["value_losses"]), _p(out["entropies"]),
                                        _p(out["total_losses"]), _p(out["ratio"]),
                                        _p(out["dlogits"]), _p(out["dvalues"]), None))
        res = {k: d.numpy() for k, d in out.items()}
        res["loss"] = float(loss.numpy()[0])
        return res


# ---- configs/*.yaml: Config / load_config — src/bin/train.cc:33-60, 108-136 ---------------------
class Config(C.Structure):
    """The reference's flat YAML schema (alegpu_config); load with Config.load(path)."""
    _fields_ = [("total_environments", C.c_int64), ("hidden_size", C.c_int64), ("horizon", C.c_int64),
                ("max_steps", C.c_int64), ("frame_stack", C.c_int64), ("learning_rate", C.c_double),
                ("clip_param", C.c_float), ("value_loss_coef", C.c_float), ("entropy_coef", C.c_float),
                ("num_epochs", C.c_int64), ("mini_batch_size", C.c_int64),
                ("num_mini_batches", C.c_int64), ("gae_discount", C.c_float),
                ("gae_lambda", C.c_float), ("max_gradient_norm", C.c_float),
                ("num_rollouts", C.c_int64), ("num_workers", C.c_int64),
                ("worker_batch_size", C.c_int64), ("frame_skip", C.c_int64), ("max_return", C.c_float),
                ("record_observation", C.c_int), ("record_video", C.c_int), ("cuda_graph", C.c_int),
                ("deterministic", C.c_int), ("action_size", C.c_int64),
                ("normalize_advantages", C.c_int), ("clip_value_loss", C.c_int), ("shuffle", C.c_int),
                ("num_gpus", C.c_int64)]

    @classmethod
    def defaults(cls):
        c = cls()
        _check(lib().alegpu_config_defaults(C.byref(c)))
        return c

    @classmethod
    def load(cls, path, validate=True):
        c = cls()
        _check(lib().alegpu_config_load(str(path).encode(), C.byref(c)))
        if validate:
            c.validate()
        return c

    @classmethod
    def parse(cls, text, validate=True):
        c = cls()
        _check(lib().alegpu_config_parse(text.encode(), C.byref(c)))
        if validate:
            c.validate()
        return c

    def validate(self):
        _check(lib().alegpu_config_validate(C.byref(self)))

    def hyperparameters(self):
        """prepare_hyperparameters — src/bin/train.cc:285-291."""
        return Hyperparameters(self.clip_param, self.value_loss_coef, self.entropy_coef,
                               self.max_gradient_norm)

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# ---- structs ----------------------------------------------------------------------------------
class _LearnerConfig(C.Structure):
    _fields_ = [("hidden_size", C.c_int64), ("action_size", C.c_int64),
                ("max_minibatch", C.c_int64), ("math", C.c_int), ("learning_rate", C.c_double),
                ("device", C.c_int), ("world_size", C.c_int), ("rank", C.c_int),
                ("nccl_unique_id", C.c_void_p)]


class Hyperparameters(C.Structure):
    """ai::ppo::train::Hyperparameters — src/ai/ppo/train.h:27-32."""
    _fields_ = [("clip_param", C.c_float), ("value_loss_coef", C.c_float),
                ("entropy_coef", C.c_float), ("max_gradient_norm", C.c_float)]


class _Batch(C.Structure):
    _fields_ = [("observations", C.c_void_p), ("actions", C.c_void_p),
                ("log_probabilities", C.c_void_p), ("advantages", C.c_void_p),
                ("returns", C.c_void_p), ("masks", C.c_void_p), ("size", C.c_int64),
                ("values", C.c_void_p)]


class _Metrics(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in
                ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
                 "advantages", "returns", "masks", "clipped_gradients")]


class _RolloutView(C.Structure):
    _fields_ = [("E", C.c_int64), ("T", C.c_int64), ("A", C.c_int64)] + [
        (k, C.c_void_p) for k in
        ("observations", "carry", "actions", "rewards", "values", "advantages", "returns",
         "logits", "log_probabilities", "terminals", "truncations", "episode_starts", "masks")]


class Batch:
    """ai::ppo::train::Batch — src/ai/ppo/train.h:34-57 (device arrays, N rows)."""

    def __init__(self, observations, actions, log_probabilities, advantages, returns, masks,
                 values=None):
        self.values = _dev(values, np.float32) if values is not None else None  # old values (clip_value_loss)
        self.observations = _dev(observations, np.uint8)
        self.actions = _dev(actions, np.int64)
        self.log_probabilities = _dev(log_probabilities, np.float32)
        self.advantages = _dev(advantages, np.float32)
        self.returns = _dev(returns, np.float32)
        self.masks = _dev(masks, np.uint8)
        self.size = int(self.actions.shape[0])

    def slice(self, start, end):
        n = end - start
        A = self.log_probabilities.shape[-1]
        return Batch(self.observations.view(start * STACK_BYTES, (n, 4, 84, 84)),
                     self.actions.view(start, (n,)),
                     self.log_probabilities.view(start * A, (n, A)),
                     self.advantages.view(start, (n,)), self.returns.view(start, (n,)),
                     self.masks.view(start, (n,)),
                     self.values.view(start, (n,)) if self.values is not None else None)

    def _c(self):
        return _Batch(self.observations.ptr, self.actions.ptr, self.log_probabilities.ptr,
                      self.advantages.ptr, self.returns.ptr, self.masks.ptr, self.size,
                      self.values.ptr if self.values is not None else None)


class Metrics:
    """ai::ppo::train::Metrics — src/ai/ppo/train.h:64-109 ([epochs,nmb,mb] device arrays)."""
    FIELDS = ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
              "advantages", "returns")

    def __init__(self, num_epochs, num_mini_batches, mini_batch_size):
        shape = (num_epochs, num_mini_batches, mini_batch_size)
        for k in self.FIELDS:
            setattr(self, k, DeviceArray.zeros(shape, np.float32))
        self.masks = DeviceArray.zeros(shape, np.uint8)
        self.clipped_gradients = DeviceArray.zeros((num_epochs, num_mini_batches), np.float32)

    def _c(self):
        return _Metrics(*[getattr(self, k).ptr for k in
                          ("loss", "clipped_losses", "value_losses", "entropies", "ratio",
                           "total_losses", "advantages", "returns", "masks",
                           "clipped_gradients")])

    def numpy(self):
        return {k: getattr(self, k).numpy() for k in self.FIELDS + ("masks", "clipped_gradients")}

    def summary(self, stream=None):
        """The scalars of log_data (src/bin/train.cc:180-190) reduced on the device and fetched in
        one copy (alegpu_metrics_summarize)."""
        E, K, mb = self.loss.shape
        dev = DeviceArray.zeros(8, np.float32)
        cm = self._c()
        _check(lib().alegpu_metrics_summarize(C.byref(cm), C.c_int64(E), C.c_int64(K),
                                              C.c_int64(mb), _p(dev), None, C.c_void_p(stream)))
        synchronize(stream)
        v = dev.numpy()
        return dict(zip(("mean_clipped_gradient", "mean_loss", "mean_clipped_loss",
                         "mean_value_loss", "mean_entropy", "mean_ratio", "mean_total_loss",
                         "unmasked"), (float(x) for x in v)))

    def masked(self, field, stream=None):
        """gather(metrics.<field>, metrics.masks) of src/bin/train.cc:149-155 (alegpu_masked_select)."""
        x = getattr(self, field)
        n = int(np.prod(x.shape))
        out, cnt = DeviceArray.zeros(max(n, 1), np.float32), DeviceArray.zeros(1, np.int64)
        _check(lib().alegpu_masked_select(_p(x), _p(self.masks), C.c_int64(n), _p(out), _p(cnt),
                                          C.c_void_p(stream)))
        synchronize(stream)
        return out.numpy()[:int(cnt.numpy()[0])]


class Buffer:
    """ai::buffer::Buffer — src/ai/buffer.h:27-58."""

    def __init__(self, total_environments, capacity, action_size=4, device=0):
        h = C.c_void_p()
        _check(lib().alegpu_rollout_create(C.c_int64(total_environments), C.c_int64(capacity),
                                           C.c_int64(action_size), C.c_int(device), C.byref(h)))
        self.h = h
        v = _RolloutView()
        _check(lib().alegpu_rollout_get_view(self.h, C.byref(v)))
        E, T, A = total_environments, capacity, action_size
        self.E, self.T, self.A = E, T, A
        mk = lambda ptr, shape, dt: DeviceArray(shape, dt, ptr, self)
        self.observations = mk(v.observations, (E, T, 4, 84, 84), np.uint8)
        self.carry = mk(v.carry, (2, E, 4, 84, 84), np.uint8)
        self.actions = mk(v.actions, (E, T), np.int64)
        self.rewards = mk(v.rewards, (E, T), np.float32)
        self.values = mk(v.values, (E, T), np.float32)
        self.advantages = mk(v.advantages, (E, T), np.float32)
        self.returns = mk(v.returns, (E, T), np.float32)
        self.logits = mk(v.logits, (E, T, A), np.float32)
        self.log_probabilities = mk(v.log_probabilities, (E, T, A), np.float32)
        self.terminals = mk(v.terminals, (E, T), np.uint8)
        self.truncations = mk(v.truncations, (E, T), np.uint8)
        self.episode_starts = mk(v.episode_starts, (E, T), np.uint8)
        self.masks = mk(v.masks, (E, T), np.uint8)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.alegpu_rollout_destroy(self.h)
            self.h = None

    def add(self, observations, actions, rewards, terminals, truncations, episode_starts, logits,
            values, stream=None):
        """Buffer::add — src/ai/buffer.cc:41-56 (per-step [E,...] arrays)."""
        o = None if observations is None else _dev(observations, np.uint8)
        args = (o, _dev(actions, np.int64), _dev(rewards, np.float32), _dev(terminals, np.uint8),
                _dev(truncations, np.uint8), _dev(episode_starts, np.uint8),
                _dev(logits, np.float32), _dev(values, np.float32))
        _check(lib().alegpu_rollout_add(self.h, *[_p(x) for x in args], C.c_void_p(stream)))
        synchronize(stream)  # staged temporaries must outlive the copy

    def get(self, next_values, discount, lambda_, check=True, stream=None):
        """Buffer::get + prepare_batch -> flat train Batch (views of the storage)."""
        nv = _dev(next_values, np.float32)
        b = _Batch()
        _check(lib().alegpu_rollout_get(self.h, _p(nv), C.c_float(discount), C.c_float(lambda_),
                                        C.byref(b), 0 if check else 1, C.c_void_p(stream)))
        synchronize(stream)
        N = self.E * self.T
        return Batch(self.observations.view(0, (N, 4, 84, 84)), self.actions.view(0, (N,)),
                     self.log_probabilities.view(0, (N, self.A)), self.advantages.view(0, (N,)),
                     self.returns.view(0, (N,)), self.masks.view(0, (N,)),
                     self.values.view(0, (N,)))


class _TrainOptions(C.Structure):
    _fields_ = [("indices", C.c_void_p), ("normalize_advantages", C.c_int),
                ("clip_value_loss", C.c_int)]


class Learner:
    """Network (src/bin/train.cc:230-269) + Adam (:360-362) + ai::ppo::train::train."""

    def __init__(self, hidden_size=512, action_size=4, max_minibatch=256, math=MATH_FP32,
                 learning_rate=2.5e-4, device=0, world_size=1, rank=0, nccl_unique_id=None):
        self._uid = None
        if nccl_unique_id is not None:
            self._uid = C.create_string_buffer(bytes(nccl_unique_id), 128)
        cfg = _LearnerConfig(hidden_size, action_size, max_minibatch, math, learning_rate, device,
                             world_size, rank,
                             C.cast(self._uid, C.c_void_p) if self._uid is not None else None)
        h = C.c_void_p()
        _check(lib().alegpu_learner_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.hidden_size, self.action_size, self.max_minibatch = hidden_size, action_size, max_minibatch
        self.P = int(lib().alegpu_learner_param_count(self.h))
        self.stream = lib().alegpu_learner_stream(self.h)

    def comm_mode(self):
        """'single', 'fused-nvlink' or 'nccl' (alegpu_learner_comm_mode)."""
        return lib().alegpu_learner_comm_mode(self.h).decode()

    def close(self):
        """Destroy the learner now (NCCL communicator included); with world_size > 1 every rank
        must call this at the same point of the program."""
        if getattr(self, "h", None) and _lib is not None:
            _lib.alegpu_learner_destroy(self.h)
            self.h = None

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.alegpu_learner_destroy(self.h)
            self.h = None

    def set_params(self, params):
        p = np.ascontiguousarray(params, np.float32)
        assert p.size == self.P
        _check(lib().alegpu_learner_set_params(self.h, p.ctypes.data_as(C.c_void_p)))

    def get_params(self):
        p = np.zeros(self.P, np.float32)
        _check(lib().alegpu_learner_get_params(self.h, p.ctypes.data_as(C.c_void_p)))
        return p

    def get_adam(self):
        m, v = np.zeros(self.P, np.float32), np.zeros(self.P, np.float32)
        st = C.c_int64(0)
        _check(lib().alegpu_learner_get_adam(self.h, m.ctypes.data_as(C.c_void_p),
                                             v.ctypes.data_as(C.c_void_p), C.byref(st)))
        return m, v, int(st.value)

    def set_adam(self, exp_avg, exp_avg_sq, step):
        m = np.ascontiguousarray(exp_avg, np.float32)
        v = np.ascontiguousarray(exp_avg_sq, np.float32)
        assert m.size == self.P and v.size == self.P
        _check(lib().alegpu_learner_set_adam(self.h, m.ctypes.data_as(C.c_void_p),
                                             v.ctypes.data_as(C.c_void_p), C.c_int64(step)))

    def save(self, path):
        """Checkpoint: parameters (reference order / OIHW), Adam moments, step -> one .npz file."""
        m, v, step = self.get_adam()
        np.savez(path, params=self.get_params(), exp_avg=m, exp_avg_sq=v, step=np.int64(step),
                 hidden_size=self.hidden_size, action_size=self.action_size)

    def load(self, path):
        z = np.load(path)
        assert int(z["hidden_size"]) == self.hidden_size and int(z["action_size"]) == self.action_size
        self.set_params(z["params"])
        self.set_adam(z["exp_avg"], z["exp_avg_sq"], int(z["step"]))

    def set_lr(self, lr):
        _check(lib().alegpu_learner_set_lr(self.h, C.c_double(lr)))

    def forward(self, observations):
        o = _dev(observations, np.uint8)
        n = o.shape[0]
        lg = DeviceArray.zeros((n, self.action_size), np.float32)
        v = DeviceArray.zeros(n, np.float32)
        _check(lib().alegpu_learner_forward(self.h, _p(o), C.c_int64(n), _p(lg), _p(v)))
        synchronize(self.stream)
        return lg.numpy(), v.numpy()

    def act(self, observations, uniforms=None, seed=0, step=0, buffer=None, t=None):
        """The action selector of src/bin/train.cc:367-379 (alegpu_learner_act): returns
        (actions [E] i64, logits [E,A], values [E]).  With buffer/t the three outputs are ALSO
        written into column t of the rollout storage (Buffer::add's job, buffer.cc:41-56)."""
        o = _dev(observations, np.uint8)
        E, A = o.shape[0], self.action_size
        u = _dev(uniforms, np.float32) if uniforms is not None else None
        dense = DeviceArray.zeros(E, np.int64)
        if buffer is not None:
            T = buffer.T
            lg, v, ac = buffer.logits, buffer.values, buffer.actions
            lp, vp, ap = lg.ptr + 4 * t * A, v.ptr + 4 * t, ac.ptr + 8 * t
            ls, vs, as_ = T * A, T, T
        else:
            lg, v = DeviceArray.zeros((E, A), np.float32), DeviceArray.zeros(E, np.float32)
            lp, vp, ap, ls, vs, as_ = lg.ptr, v.ptr, None, A, 1, 0
        _check(lib().alegpu_learner_act(self.h, _p(o), C.c_int64(E), _p(u) if u is not None else None,
                                        C.c_uint64(seed), C.c_uint64(step), C.c_void_p(lp),
                                        C.c_int64(ls), C.c_void_p(vp), C.c_int64(vs),
                                        C.c_void_p(ap), C.c_int64(as_), _p(dense), None))
        synchronize(self.stream)
        if buffer is not None:
            return (dense.numpy(), lg.numpy().reshape(E, T, A)[:, t], v.numpy().reshape(E, T)[:, t])
        return dense.numpy(), lg.numpy(), v.numpy()

    def activations(self, n):
        c1 = np.zeros((n, 32, 20, 20), np.float32)
        c2 = np.zeros((n, 64, 9, 9), np.float32)
        c3 = np.zeros((n, 64, 7, 7), np.float32)
        fc = np.zeros((n, self.hidden_size), np.float32)
        _check(lib().alegpu_learner_get_activations(
            self.h, C.c_int64(n), *[x.ctypes.data_as(C.c_void_p) for x in (c1, c2, c3, fc)]))
        return c1, c2, c3, fc

    def grads(self, batch, hp):
        g = np.zeros(self.P, np.float32)
        loss = C.c_float(0)
        n = batch.size
        lg = DeviceArray.zeros((n, self.action_size), np.float32)
        v = DeviceArray.zeros(n, np.float32)
        cb = batch._c()
        _check(lib().alegpu_learner_grads(self.h, C.byref(cb), C.byref(hp),
                                          g.ctypes.data_as(C.c_void_p), C.byref(loss), _p(lg),
                                          _p(v)))
        return g, float(loss.value), lg.numpy(), v.numpy()

    def apply_grads(self, grads, max_norm):
        g = np.ascontiguousarray(grads, np.float32)
        tn = C.c_float(0)
        _check(lib().alegpu_learner_apply_grads(self.h, g.ctypes.data_as(C.c_void_p),
                                                C.c_double(max_norm), C.byref(tn)))
        return float(tn.value)

    def train(self, batch, metrics, num_epochs, num_mini_batches, hp, cuda_graph=False,
              sync=True, indices=None, normalize_advantages=False, clip_value_loss=False):
        """ai::ppo::train::train — src/ai/ppo/train.h:133-157.  indices ([epochs*N] int64 row
        order), normalize_advantages and clip_value_loss (needs batch.values) are the opt-in extras
        of alegpu_train_options."""
        cb = batch._c()
        cm = metrics._c() if metrics is not None else None
        self._idx = _dev(indices, np.int64) if indices is not None else None
        opt = _TrainOptions(self._idx.ptr if self._idx is not None else None,
                            1 if normalize_advantages else 0, 1 if clip_value_loss else 0)
        _check(lib().alegpu_learner_train_ex(self.h, C.byref(cb),
                                             C.byref(cm) if cm is not None else None,
                                             C.c_int64(num_epochs), C.c_int64(num_mini_batches),
                                             C.byref(hp), C.byref(opt), 1 if cuda_graph else 0))
        if sync:
            synchronize(self.stream)

    def last_launch_count(self):
        return int(lib().alegpu_learner_last_launch_count(self.h))


# ---- env -> GPU mapping of the data-parallel path -------------------------------------
def shard_rows(N, num_mini_batches, world_size, rank):
    """Rows of the flat [N] batch that rank `rank` owns so that ITS minibatch k is its share of the
    reference's minibatch k = rows [k*mb, (k+1)*mb) (src/ai/ppo/train.h:149-151): for every k the
    slice [k*mb + rank*mb/G, k*mb + (rank+1)*mb/G).  With this mapping G GPUs train on exactly the
    minibatches one GPU would (up to fp32 summation order)."""
    mb = N // num_mini_batches
    assert N % num_mini_batches == 0 and mb % world_size == 0, "mb must divide over the ranks"
    per = mb // world_size
    return np.concatenate([np.arange(k * mb + rank * per, k * mb + (rank + 1) * per)
                           for k in range(num_mini_batches)])


def shard_envs(E, T, num_mini_batches, world_size, rank):
    """The same mapping in env units (rows are env-major, row = e*T + t): valid when the reference's
    minibatches are whole envs (num_mini_batches * world_size divides E; v1: 4096/(16*8) = 32, sweep:
    4096/(32*8) = 16): rank g owns envs [k*E/nmb + g*E/(nmb*G), k*E/nmb + (g+1)*E/(nmb*G)) for every k."""
    assert E % (num_mini_batches * world_size) == 0, "nmb * world_size must divide E"
    per_mb, per = E // num_mini_batches, E // (num_mini_batches * world_size)
    return np.concatenate([np.arange(k * per_mb + rank * per, k * per_mb + (rank + 1) * per)
                           for k in range(num_mini_batches)])


def selftest_tc_gemm(A, B, M, N, R, a_mn=False, b_mn=False, bn=64, splits=1, use_async=False,
                     tf32=False):
    """C[M,N] = A . B^T through the tcgen05/TMEM engine (alegpu_selftest_tc_gemm).  A is
    [M,R] (a_mn False) or [R,M] (a_mn True) fp32; likewise B.  Returns [z,M,N] partial
    products (z == 1 unless splits > 1)."""
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    z = 1
    if splits > 1:
        rps = -(-(-(-R // splits)) // 64) * 64
        z = -(-R // rps)
    dA, dB = DeviceArray.from_numpy(A), DeviceArray.from_numpy(B)
    dC = DeviceArray.zeros((z, M, N), np.float32)
    fn = lib().alegpu_selftest_tc_gemm_async if use_async else lib().alegpu_selftest_tc_gemm
    if tf32:  # kind::tf32: fp32 inputs read as they are (upper 19 bits) by the cp.async producers
        fn = lib().alegpu_selftest_tc_gemm_tf32
    _check(fn(_p(dA), _p(dB), _p(dC), C.c_int64(M), C.c_int64(N), C.c_int64(R),
              C.c_int(int(a_mn)), C.c_int(int(b_mn)), C.c_int(bn), C.c_int(splits), None))
    synchronize()
    return dC.numpy()


# ---- ai::rollout::Rollout::rollout() — src/ai/rollout.cc:198-278 ------------------------------------
class _RolloutLog(C.Structure):
    _fields_ = [("steps", C.c_int64), ("episodes", C.c_int64), ("finished_episodes", C.c_int64),
                ("finished_games", C.c_int64), ("mean_episode_return", C.c_float),
                ("mean_episode_length", C.c_float), ("mean_game_return", C.c_float),
                ("mean_game_length", C.c_float)]


ENV_STEP_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.POINTER(C.c_uint8),
                          C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_uint8),
                          C.POINTER(C.c_uint8), C.POINTER(C.c_uint8))


class Collector:
    """Rollout collection (alegpu_collector_*): CPU emulator threads behind a step callback (None =
    the built-in Breakout-shaped frame source), pinned frame ring, per-step hand-off."""

    def __init__(self, total_environments, horizon, action_size=4, num_workers=1, step_fn=None,
                 seed=0, device=0):
        self._fn = ENV_STEP_FN(step_fn) if step_fn is not None else None
        h = C.c_void_p()
        _check(lib().alegpu_collector_create(
            C.c_int64(total_environments), C.c_int64(horizon), C.c_int64(action_size),
            C.c_int(num_workers), self._fn if self._fn is not None else C.cast(None, ENV_STEP_FN),
            None, C.c_uint64(seed), C.c_int(device), C.byref(h)))
        self.h = h
        self.E, self.T, self.A = total_environments, horizon, action_size

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.alegpu_collector_destroy(self.h)
            self.h = None

    def rollout(self, buffer, learner, gamma, lambda_):
        """Returns (Batch of [N] views, log dict); synchronises the learner's stream."""
        b, log = _Batch(), _RolloutLog()
        _check(lib().alegpu_collector_rollout(self.h, buffer.h, learner.h, C.c_float(gamma),
                                              C.c_float(lambda_), C.byref(b), C.byref(log)))
        synchronize(learner.stream)
        N = self.E * self.T
        batch = Batch(buffer.observations.view(0, (N, 4, 84, 84)), buffer.actions.view(0, (N,)),
                      buffer.log_probabilities.view(0, (N, self.A)),
                      buffer.advantages.view(0, (N,)), buffer.returns.view(0, (N,)),
                      buffer.masks.view(0, (N,)))
        return batch, {k: getattr(log, k) for k, _ in _RolloutLog._fields_}

    def frames(self):
        """The pinned host frame ring of the last rollout as a numpy view [T,E,2,210,160,3]."""
        lib().alegpu_collector_frames.restype = C.POINTER(C.c_uint8)
        p = lib().alegpu_collector_frames(self.h)
        return np.ctypeslib.as_array(p, shape=(self.T, self.E, 2, 210, 160, 3))
