"""ctypes binding of libmatcha_b200.so (the C ABI declared in include/matcha_b200.h).

There is no CPU or PyTorch fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmatcha_b200.so")

F32, BF16, F16 = 0, 1, 2
DTYPES = {"fp32": F32, "float32": F32, "bf16": BF16, "bfloat16": BF16, "fp16": F16, "float16": F16, "half": F16}

EXPORTS = [
    "mb200_last_error", "mb200_version", "mb200_plan_create", "mb200_plan_destroy", "mb200_load_estimator",
    "mb200_load_generator", "mb200_load_text_encoder", "mb200_text_encoder_forward", "mb200_euler_schedule", "mb200_estimator_forward", "mb200_cfm_decode",
    "mb200_hifigan_forward", "mb200_synthesise", "mb200_synthesise_host", "mb200_synthesise_host_submit", "mb200_synthesise_host_wait", "mb200_launch_count",
    "mb200_workspace_bytes", "mb200_compute_loss", "mb200_set_train_dropout", "mb200_denoiser_bias", "mb200_denoise", "mb200_length_regulate", "mb200_maximum_path", "mb200_maximum_path_workspace_bytes", "mb200_maximum_path_static", "mb200_set_profiling", "mb200_get_profile", "mb200_test_conv1d", "mb200_test_attention", "mb200_debug_probe",
]


class Cfg(ctypes.Structure):
    _fields_ = [("dtype", ctypes.c_int), ("c_in", ctypes.c_int), ("force_simt", ctypes.c_int),
                ("use_graph", ctypes.c_int)]


class Tensor(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("data", ctypes.c_void_p), ("ndim", ctypes.c_int),
                ("shape", ctypes.c_longlong * 4)]


class Param(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("param", ctypes.c_void_p), ("grad", ctypes.c_void_p)]


_lib = None
_lock = threading.Lock()


def load():
    """Loads the shared library (once).  Raises if it has not been built: the product path has no fallback."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m matcha_b200.build` (nvcc, sm_100a). "
                "matcha_b200 has no CPU / PyTorch fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        lib.mb200_last_error.restype = ctypes.c_char_p
        lib.mb200_version.restype = ctypes.c_char_p
        lib.mb200_plan_create.argtypes = [ctypes.POINTER(Cfg), ctypes.POINTER(vp)]
        lib.mb200_plan_destroy.argtypes = [vp]
        lib.mb200_plan_destroy.restype = None
        lib.mb200_load_estimator.argtypes = [vp, ctypes.POINTER(Tensor), i, ctypes.c_char_p]
        lib.mb200_load_generator.argtypes = [vp, ctypes.POINTER(Tensor), i]
        lib.mb200_load_text_encoder.argtypes = [vp, ctypes.POINTER(Tensor), i, ctypes.c_char_p, i]
        lib.mb200_text_encoder_forward.argtypes = [vp, vp, vp, vp, i, i, vp, vp, vp, vp]
        lib.mb200_euler_schedule.argtypes = [i, ctypes.POINTER(f), ctypes.POINTER(f)]
        lib.mb200_estimator_forward.argtypes = [vp, vp, vp, vp, vp, i, vp, i, i, vp, vp]
        lib.mb200_cfm_decode.argtypes = [vp, vp, vp, vp, vp, i, i, i, vp, vp]
        lib.mb200_hifigan_forward.argtypes = [vp, vp, i, i, vp, vp]
        lib.mb200_synthesise.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, f, f, vp, vp, vp, vp]
        lib.mb200_synthesise_host.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, f, f, vp, vp]
        lib.mb200_synthesise_host_submit.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, f, f, vp, vp, i]
        lib.mb200_synthesise_host_wait.argtypes = [vp, i]
        lib.mb200_compute_loss.argtypes = [vp, ctypes.POINTER(Param), i, ctypes.c_char_p, vp, vp, vp, vp, vp, vp, i, i, f, i,
                                           vp, vp, vp, vp, vp, vp]
        lib.mb200_set_train_dropout.argtypes = [vp, f, ctypes.c_ulonglong]
        lib.mb200_denoiser_bias.argtypes = [vp, vp, i, vp, vp]
        lib.mb200_denoise.argtypes = [vp, vp, i, i, vp, f, vp, vp]
        lib.mb200_length_regulate.argtypes = [vp, vp, vp, vp, i, i, i, i, vp, vp, vp, vp, vp]
        lib.mb200_maximum_path.argtypes = [vp, vp, i, i, i, vp, vp, vp]
        lib.mb200_maximum_path_static.argtypes = [vp, vp, i, i, i, vp, vp]
        lib.mb200_maximum_path_workspace_bytes.argtypes = [i, i, i]
        lib.mb200_maximum_path_workspace_bytes.restype = ctypes.c_longlong
        lib.mb200_launch_count.argtypes = [vp]
        lib.mb200_launch_count.restype = ctypes.c_longlong
        lib.mb200_workspace_bytes.argtypes = [vp]
        lib.mb200_workspace_bytes.restype = ctypes.c_longlong
        lib.mb200_set_profiling.argtypes = [vp, i]
        lib.mb200_get_profile.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
        lib.mb200_test_conv1d.argtypes = [vp, vp, vp, vp, i, i, i, i, i, i, i, i, vp, vp]
        lib.mb200_test_attention.argtypes = [vp, vp, vp, i, i, vp, vp]
        lib.mb200_debug_probe.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong), i]
        _lib = lib
        return lib


def check(rc):
    if rc != 0:
        raise RuntimeError("matcha_b200: " + load().mb200_last_error().decode("utf-8", "replace"))


def euler_schedule(n_steps):
    """(t, dt) per step exactly as the CUDA path uses them (host-only call; no GPU needed)."""
    lib = load()
    ts = (ctypes.c_float * n_steps)()
    dts = (ctypes.c_float * n_steps)()
    check(lib.mb200_euler_schedule(n_steps, ts, dts))
    return list(ts), list(dts)


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _dev_f32(t, device):
    if t is None:
        return None
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def length_regulate(mu_x, w, x_mask, y_lengths, t_pad, want_attn=True):
    """CUDA length regulation (matcha_tts.py:118-131): mu_x (B, C, Lx), w (B, Lx) frames per token, x_mask (B, Lx),
    y_lengths (B,) int64 -> mu_y (B, C, t_pad), attn (B, Lx, t_pad) or None, y_mask (B, t_pad).  No plan needed."""
    lib = load()
    if not mu_x.is_cuda:
        raise RuntimeError("matcha_b200.length_regulate runs on CUDA tensors only (no CPU fallback)")
    dev = mu_x.device
    mu_x, w, x_mask = _dev_f32(mu_x, dev), _dev_f32(w, dev), _dev_f32(x_mask, dev)
    b, c, lx = mu_x.shape
    if tuple(w.shape) != (b, lx) or tuple(x_mask.shape) != (b, lx):
        raise ValueError("w and x_mask must be (B, Lx)")
    yl = y_lengths.to(device=dev, dtype=torch.int64).contiguous()
    if tuple(yl.shape) != (b,):
        raise ValueError("y_lengths must be (B,)")
    t_pad = int(t_pad)
    cum = torch.empty(b, lx, device=dev, dtype=torch.float32)
    mu_y = torch.empty(b, c, t_pad, device=dev, dtype=torch.float32)
    attn = torch.empty(b, lx, t_pad, device=dev, dtype=torch.float32) if want_attn else None
    y_mask = torch.empty(b, t_pad, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        check(lib.mb200_length_regulate(_ptr(mu_x), _ptr(w), _ptr(x_mask), _ptr(yl), b, c, lx, t_pad, _ptr(cum), _ptr(mu_y),
                                        _ptr(attn), _ptr(y_mask), _stream()))
    return mu_y, attn, y_mask


def maximum_path(value, mask):
    """CUDA monotonic alignment search: value, mask (B, Tx, Ty) -> 0/1 path (B, Tx, Ty) in value's dtype
    (monotonic_align/__init__.py:7-22).  No plan needed, no host round trip."""
    lib = load()
    if not value.is_cuda:
        raise RuntimeError("matcha_b200.maximum_path runs on CUDA tensors only (no CPU fallback)")
    dev = value.device
    if value.dim() != 3 or tuple(mask.shape) != tuple(value.shape):
        raise ValueError("value and mask must both be (B, Tx, Ty)")
    v, m = _dev_f32(value, dev), _dev_f32(mask, dev)
    b, tx, ty = v.shape
    path = torch.empty(b, tx, ty, device=dev, dtype=torch.float32)
    ws = torch.empty(max(1, int(lib.mb200_maximum_path_workspace_bytes(b, tx, ty))), device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        check(lib.mb200_maximum_path(_ptr(v), _ptr(m), b, tx, ty, _ptr(path), _ptr(ws), _stream()))
    return path.to(value.dtype)


MAS_STATIC_SCRATCH_BYTES = 4 * 1024 * 1024   # MB200_MAS_STATIC_SCRATCH_BYTES (include/matcha_b200.h)


def maximum_path_static(value, mask, path):
    """`mb200_maximum_path_static`: the same search with the kernel's scratch in a static device array inside the library.
    Nothing is allocated here or below: value, mask (B, Tx, Ty) and the output `path` are the caller's contiguous fp32
    CUDA tensors.  Raises when the shape needs more than MAS_STATIC_SCRATCH_BYTES of scratch.  Calls that overlap on one
    GPU race on the shared array (no counterpart in the reference)."""
    lib = load()
    for t in (value, mask, path):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise ValueError("maximum_path_static takes contiguous fp32 CUDA tensors (it converts and allocates nothing)")
    if value.dim() != 3 or tuple(mask.shape) != tuple(value.shape) or tuple(path.shape) != tuple(value.shape):
        raise ValueError("value, mask and path must all be (B, Tx, Ty)")
    b, tx, ty = value.shape
    with torch.cuda.device(value.device):
        check(lib.mb200_maximum_path_static(_ptr(value), _ptr(mask), b, tx, ty, _ptr(path), _stream()))
    return path


class Plan:
    """Owns one mb200_plan: packed weights, workspaces, CUDA graphs.  One per (precision, c_in, device)."""

    def __init__(self, precision="fp16", c_in=160, force_simt=False, use_graph=True, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("matcha_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.precision = precision
        self.c_in = c_in
        cfg = Cfg(DTYPES[precision], c_in, int(force_simt), int(use_graph))
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(self.lib.mb200_plan_create(ctypes.byref(cfg), ctypes.byref(handle)))
        self.handle = handle
        self._keep = []

    def __del__(self):
        h, self.handle = getattr(self, "handle", None), None
        if h:
            try:
                self.lib.mb200_plan_destroy(h)
            except Exception:
                pass

    # ---- weights
    def _tensor_array(self, state_dict):
        items = [(k, v.detach().to("cpu", torch.float32).contiguous()) for k, v in state_dict.items()
                 if torch.is_tensor(v)]
        arr = (Tensor * len(items))()
        keep = []
        for j, (k, v) in enumerate(items):
            name = k.encode()
            keep.append((name, v))
            arr[j].name = name
            arr[j].data = v.data_ptr()
            arr[j].ndim = v.ndim
            for d in range(v.ndim):
                arr[j].shape[d] = v.shape[d]
        return arr, len(items), keep

    def load_estimator(self, state_dict, prefix="estimator."):
        arr, n, keep = self._tensor_array(state_dict)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_load_estimator(self.handle, arr, n, prefix.encode()))
        del keep

    def load_generator(self, state_dict):
        arr, n, keep = self._tensor_array(state_dict)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_load_generator(self.handle, arr, n))
        del keep

    def load_text_encoder(self, state_dict, n_heads=2, prefix=""):
        arr, n, keep = self._tensor_array(state_dict)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_load_text_encoder(self.handle, arr, n, prefix.encode(), int(n_heads)))
        del keep

    # ---- compute
    def text_encoder_forward(self, x, x_lengths, spks=None):
        """TextEncoder.forward on the CUDA kernels: ids (B, Lx), lengths (B,) -> mu_x (B,80,Lx), logw (B,1,Lx), x_mask (B,1,Lx)."""
        b, lx = x.shape
        ids = x.detach().to(device=self.device, dtype=torch.int64).contiguous()
        lens = x_lengths.detach().to(device=self.device, dtype=torch.int64).contiguous()
        spks = _dev_f32(spks, self.device)
        mu_x = torch.empty(b, 80, lx, device=self.device, dtype=torch.float32)
        logw = torch.empty(b, 1, lx, device=self.device, dtype=torch.float32)
        x_mask = torch.empty(b, 1, lx, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_text_encoder_forward(self.handle, _ptr(ids), _ptr(lens), _ptr(spks), b, lx, _ptr(mu_x),
                                                      _ptr(logw), _ptr(x_mask), _stream()))
        return mu_x, logw, x_mask

    def estimator_forward(self, x, mask, mu, t, spks=None):
        b, c, tt = x.shape
        x, mask, mu, spks = (_dev_f32(v, self.device) for v in (x, mask, mu, spks))
        t = _dev_f32(torch.as_tensor(t), self.device).reshape(-1)
        batched = t.numel() > 1
        if batched and t.numel() != b:
            raise ValueError("t must be a scalar or have one value per batch element")
        out = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_estimator_forward(self.handle, _ptr(x), _ptr(mask), _ptr(mu), _ptr(t), int(batched),
                                                   _ptr(spks), b, tt, _ptr(out), _stream()))
        return out

    def cfm_decode(self, mu, mask, z, n_steps, spks=None):
        b, c, tt = mu.shape
        mu, mask, z, spks = (_dev_f32(v, self.device) for v in (mu, mask, z, spks))
        out = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_cfm_decode(self.handle, _ptr(mu), _ptr(mask), _ptr(z), _ptr(spks), b, tt, int(n_steps),
                                            _ptr(out), _stream()))
        return out

    def hifigan_forward(self, mel):
        b, c, l = mel.shape
        if c != 80:
            raise ValueError("HiFi-GAN v1 expects 80 mel channels")
        mel = _dev_f32(mel, self.device)
        wav = torch.empty(b, 1, 256 * l, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_hifigan_forward(self.handle, _ptr(mel), b, l, _ptr(wav), _stream()))
        return wav

    def synthesise(self, mu, mask, z, n_steps, length, mel_mean, mel_std, spks=None):
        """decode -> crop to `length` frames -> denormalise -> vocode.  Returns (decoded, mel, wav)."""
        b, c, tt = mu.shape
        mu, mask, z, spks = (_dev_f32(v, self.device) for v in (mu, mask, z, spks))
        dec = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32)
        mel = torch.empty(b, 80, length, device=self.device, dtype=torch.float32)
        wav = torch.empty(b, 1, 256 * length, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_synthesise(self.handle, _ptr(mu), _ptr(mask), _ptr(z), _ptr(spks), b, tt, int(n_steps),
                                            int(length), float(mel_mean), float(mel_std), _ptr(dec), _ptr(mel),
                                            _ptr(wav), _stream()))
        return dec, mel, wav

    def synthesise_host(self, mu, mask, z, n_steps, length, mel_mean, mel_std, spks=None, mel_out=None, wav_out=None):
        """Host-buffer end-to-end call (H2D, hot path, D2H, sync).  Inputs are CPU fp32 tensors (ideally pinned)."""
        b, c, tt = mu.shape
        for v in (mu, mask, z, spks):
            if v is not None and (v.is_cuda or v.dtype != torch.float32 or not v.is_contiguous()):
                raise ValueError("synthesise_host takes contiguous fp32 CPU tensors")
        if mel_out is None:
            mel_out = torch.empty(b, 80, length, dtype=torch.float32).pin_memory()
        if wav_out is None:
            wav_out = torch.empty(b, 1, 256 * length, dtype=torch.float32).pin_memory()
        with torch.cuda.device(self.device):
            check(self.lib.mb200_synthesise_host(self.handle, _ptr(mu), _ptr(mask), _ptr(z), _ptr(spks), b, tt,
                                                 int(n_steps), int(length), float(mel_mean), float(mel_std),
                                                 _ptr(mel_out), _ptr(wav_out)))
        return mel_out, wav_out

    def synthesise_host_submit(self, mu, mask, z, n_steps, length, mel_mean, mel_std, mel_out, wav_out, spks=None, slot=0):
        """First half of synthesise_host: queue one batch (pinned fp32 CPU tensors in, pinned outputs) and return.  Pair
        with synthesise_host_wait(slot); two slots allow the copies of one batch to overlap the compute of the other."""
        b, c, tt = mu.shape
        for v in (mu, mask, z, spks, mel_out, wav_out):
            if v is not None and (v.is_cuda or v.dtype != torch.float32 or not v.is_contiguous()):
                raise ValueError("synthesise_host_submit takes contiguous fp32 CPU tensors")
        with torch.cuda.device(self.device):
            check(self.lib.mb200_synthesise_host_submit(self.handle, _ptr(mu), _ptr(mask), _ptr(z), _ptr(spks), b, tt,
                                                        int(n_steps), int(length), float(mel_mean), float(mel_std),
                                                        _ptr(mel_out), _ptr(wav_out), int(slot)))

    def synthesise_host_wait(self, slot=0):
        with torch.cuda.device(self.device):
            check(self.lib.mb200_synthesise_host_wait(self.handle, int(slot)))

    def set_train_dropout(self, p, seed=0):
        """Transformer-block dropout for the following compute_loss calls (0 = eval mode)."""
        check(self.lib.mb200_set_train_dropout(self.handle, float(p), int(seed) & 0xFFFFFFFFFFFFFFFF))

    def compute_loss(self, named_params, x1, mask, mu, z, t, spks=None, sigma_min=1e-4, backward=True, want_dmu=False,
                     want_dspks=False):
        """BASECFM.compute_loss forward (+ backward) with injected draws t (B,), z (B,80,T).

        named_params: iterable of (name, tensor) -- the estimator's parameters, CUDA fp32 contiguous, reference names
        (SURVEY.md Appendix A).  Returns dict(loss, y, v, grads {name: tensor}, dmu, dspks); gradients are fresh
        zero-initialised tensors filled by the CUDA backward pass (flow_matching.py:87-118 under autograd)."""
        b, c, tt = x1.shape
        x1, mask, mu, z, spks = (_dev_f32(v, self.device) for v in (x1, mask, mu, z, spks))
        t = _dev_f32(t, self.device).reshape(-1)
        if t.numel() != b:
            raise ValueError("t must have one value per batch element")
        named = [(k, v) for k, v in named_params]
        for k, v in named:
            if not v.is_cuda or v.dtype != torch.float32 or not v.is_contiguous():
                raise ValueError(f"parameter {k} must be a contiguous fp32 CUDA tensor (the kernels read it in place)")
        grads = {k: torch.zeros_like(v) for k, v in named} if backward else {}
        arr = (Param * len(named))()
        keep = []
        for j, (k, v) in enumerate(named):
            name = k.encode()
            keep.append(name)
            arr[j].name = name
            arr[j].param = v.data_ptr()
            arr[j].grad = grads[k].data_ptr() if backward else None
        loss = torch.zeros((), device=self.device, dtype=torch.float32)
        y = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32)
        v_out = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32)
        dmu = torch.empty(b, 80, tt, device=self.device, dtype=torch.float32) if backward and want_dmu else None
        dspks = torch.empty_like(spks) if backward and want_dspks and spks is not None else None
        with torch.cuda.device(self.device):
            check(self.lib.mb200_compute_loss(self.handle, arr, len(named), b"", _ptr(x1), _ptr(mask), _ptr(mu), _ptr(z),
                                              _ptr(t), _ptr(spks), b, tt, float(sigma_min), int(backward), _ptr(loss),
                                              _ptr(y), _ptr(v_out), _ptr(dmu), _ptr(dspks), _stream()))
        del keep
        return {"loss": loss, "y": y, "v": v_out, "grads": grads, "dmu": dmu, "dspks": dspks}

    def denoiser_bias(self, bias_audio):
        """|STFT(bias_audio (L,))|[:, 0] -> (513,)  (hifigan/denoiser.py:53-57)."""
        a = _dev_f32(bias_audio, self.device).reshape(-1)
        bias = torch.empty(513, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_denoiser_bias(self.handle, _ptr(a), a.numel(), _ptr(bias), _stream()))
        return bias

    def denoise(self, audio, bias, strength):
        """audio (B, L) -> (B, 256 * (L // 256))  (hifigan/denoiser.py:59-64)."""
        a = _dev_f32(audio, self.device)
        b, l = a.shape
        bias = _dev_f32(bias, self.device).reshape(-1)
        if bias.numel() != 513:
            raise ValueError("bias must hold 513 bins")
        out = torch.empty(b, 256 * (l // 256), device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_denoise(self.handle, _ptr(a), b, l, _ptr(bias), float(strength), _ptr(out), _stream()))
        return out

    def set_profiling(self, on):
        check(self.lib.mb200_set_profiling(self.handle, int(on)))

    def get_profile(self):
        """{(region, kind): (launches, ms, flops)} for the non-empty slots; see include/matcha_b200.h."""
        buf = (ctypes.c_double * (6 * 4 * 3))()
        check(self.lib.mb200_get_profile(self.handle, buf))
        out = {}
        for r in range(6):
            for k in range(4):
                n, ms, fl = buf[(r * 4 + k) * 3:(r * 4 + k) * 3 + 3]
                if n:
                    out[(r, k)] = (int(n), ms, fl)
        return out

    def launch_count(self):
        return int(self.lib.mb200_launch_count(self.handle))

    def workspace_bytes(self):
        return int(self.lib.mb200_workspace_bytes(self.handle))

    def debug_probe(self, n_ctas=148):
        """[n_ctas][16] cycle counters of the last instrumented kernel (developer switches MB200_*_DBG)."""
        buf = (ctypes.c_longlong * (n_ctas * 16))()
        check(self.lib.mb200_debug_probe(self.handle, buf, n_ctas))
        return [list(buf[c * 16:(c + 1) * 16]) for c in range(n_ctas)]

    # ---- unit-test hooks
    def test_conv1d(self, x, w, bias, dilation=1, stride=1, transposed=False):
        b, cin, t = x.shape
        k = w.shape[2]
        cout = w.shape[1] if transposed else w.shape[0]
        t_out = t * stride if transposed else t // stride
        x = _dev_f32(x, self.device)
        w = w.detach().to("cpu", torch.float32).contiguous()
        bias = None if bias is None else bias.detach().to("cpu", torch.float32).contiguous()
        y = torch.empty(b, cout, t_out, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_test_conv1d(self.handle, _ptr(x), _ptr(w), _ptr(bias), b, cin, cout, t, k, dilation,
                                             stride, int(transposed), _ptr(y), _stream()))
        return y

    def test_attention(self, qkv, mask):
        b, t, _ = qkv.shape
        qkv, mask = _dev_f32(qkv, self.device), _dev_f32(mask, self.device)
        out = torch.empty(b, t, 128, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            check(self.lib.mb200_test_attention(self.handle, _ptr(qkv), _ptr(mask), b, t, _ptr(out), _stream()))
        return out
