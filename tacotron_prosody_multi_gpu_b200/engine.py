"""Host-side driver of the C-ABI extension: one ``TacoEngine`` per GPU.

PyTorch is only the carrier of device memory and streams; every computation is a
``taco_*`` call into libtaco_b200.so (include/taco_b200.h).  The reference-facing
classes in ``TacotronModel/`` and ``Utils/`` are thin layers over this engine.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .hparams import num_symbols

MAX_BATCH_PER_CALL = 64  # one persistent-kernel launch covers <= 64 utterances; larger batches are chunked


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


class TacoEngine(object):
    def __init__(self, hp, weights=None, device=0, precision="fp32"):
        if not torch.cuda.is_available():
            raise _lib.TacoError("TacoEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.hp = hp
        self.device = torch.device("cuda", device)
        self.L = _lib.lib()
        self._h = C.c_void_p()
        hps = _lib.make_hparams(hp, num_symbols(hp))
        rc = self.L.taco_create(C.byref(hps), device, C.byref(self._h))
        if rc != 0:
            msg = self.L.taco_last_error(self._h).decode() if self._h else ""
            if self._h:
                self.L.taco_destroy(self._h)
                self._h = C.c_void_p()
            raise _lib.TacoError("taco_create failed: %s %s" % (_lib.ERROR_NAMES.get(rc, rc), msg))
        self.precision = "fp32"
        if weights is not None:
            self.load_weights(weights)
        if precision != "fp32":
            self.set_precision(precision)

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc, what):
        if rc != 0:
            raise _lib.TacoError("%s failed: %s: %s" % (what, _lib.ERROR_NAMES.get(rc, rc),
                                                         self.L.taco_last_error(self._h).decode()))

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "_h", None):
            self.L.taco_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_weights(self, weights):
        """``weights``: dict name -> float32 ndarray / tensor (names + TF layouts of weights.py)."""
        for name, v in weights.items():
            a = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
            a = np.ascontiguousarray(a, dtype=np.float32)
            shape = (C.c_int64 * max(a.ndim, 1))(*a.shape)
            self._check(self.L.taco_set_weight(self._h, name.encode(), a.ctypes.data_as(C.c_void_p), shape, a.ndim, 0),
                        "taco_set_weight(%s)" % name)
        self._check(self.L.taco_finalize_weights(self._h), "taco_finalize_weights")

    def set_precision(self, precision):
        code = {"fp32": _lib.TACO_PRECISION_FP32, "bf16_conv": _lib.TACO_PRECISION_BF16_CONV,
                "bf16": _lib.TACO_PRECISION_BF16}[precision]
        self._check(self.L.taco_set_precision(self._h, code), "taco_set_precision")
        self.precision = precision

    def set_encoder_ctas(self, max_ctas):
        """Grid bound of the persistent BiLSTM (include/taco_b200.h taco_set_encoder_ctas): 0 = lowest latency, 32 = fits beside a
        running cluster decoder."""
        self._check(self.L.taco_set_encoder_ctas(self._h, int(max_ctas)), "taco_set_encoder_ctas")

    @property
    def launch_count(self):
        return int(self.L.taco_launch_count(self._h))

    def stage_ms(self, stage):
        """Device milliseconds of the last run of a stage (CUDA events inside the library)."""
        ms = C.c_float()
        self._check(self.L.taco_stage_ms(self._h, int(stage), C.byref(ms)), "taco_stage_ms")
        return float(ms.value)

    # ------------------------------------------------------------------ path stages
    def encode(self, ids, lens, reference_mel=None, gst_weights=None):
        """-> memory [B, T, 768] (= ``encoder_outputs``, Tacotron.py:224)."""
        ids = ids.to(self.device, torch.int32).contiguous()
        lens = lens.to(self.device, torch.int32).contiguous()
        B, T = ids.shape
        mem_dim = 2 * self.hp.enc_lstm_hidden_size + self.hp.style_embed_depth
        memory = torch.empty(B, T, mem_dim, device=self.device, dtype=torch.float32)
        ref = gw = None
        Tm = 0
        if reference_mel is not None:
            ref = reference_mel.to(self.device, torch.float32).contiguous()
            if ref.shape[0] == 1 and B > 1:
                ref = ref.expand(B, -1, -1).contiguous()
            Tm = ref.shape[1]
        else:
            if gst_weights is None:
                raise ValueError("no reference mel: GST weights [heads, tokens] must be passed explicitly")
            gw = torch.as_tensor(gst_weights, dtype=torch.float32).to(self.device).contiguous()
        self._check(self.L.taco_encode(self._h, _ptr(ids), _ptr(lens), B, T, _ptr(ref), Tm, _ptr(gw), _ptr(memory),
                                       self._stream()), "taco_encode")
        return memory

    def decode(self, memory, lens, mel_targets=None, max_iters=None, prenet_masks=None, philox_seed=0,
               reset_state=True, logits_only=False, sync=True, single_step_at=None):
        """One persistent-kernel launch for all decoder steps.

        Returns dict(frames [B, 2*S_max, 80] raw, stop [B, 2*S_max], align_hist [S_max, B, T],
        steps (python int when ``sync`` else device tensor), S_max)."""
        memory = memory.contiguous()
        lens = lens.to(self.device, torch.int32).contiguous()
        B, T, _ = memory.shape
        r, M = self.hp.outputs_per_step, self.hp.num_mels
        tgt = None
        F_tgt = 0
        if mel_targets is not None:
            tgt = mel_targets.to(self.device, torch.float32).contiguous()
            F_tgt = tgt.shape[1]
            S_max = F_tgt // r
            if max_iters is not None:
                S_max = min(S_max, max_iters)
            if single_step_at is not None:  # `cell(inputs, state)`: step t reads targets[:, r*t - 1]
                assert F_tgt >= r * single_step_at
                S_max = 1
            mode = _lib.TACO_MODE_TEACHER
        else:
            S_max = self.hp.max_iters if max_iters is None else max_iters
            mode = _lib.TACO_MODE_FREE
        masks = None
        if prenet_masks is not None:
            masks = torch.as_tensor(prenet_masks).to(self.device, torch.uint8).contiguous()
            assert masks.shape == (S_max, B, 2, self.hp.prenet_layers[0]), masks.shape
        frames = torch.empty(B, S_max * r, M, device=self.device, dtype=torch.float32)
        stop = torch.empty(B, S_max * r, device=self.device, dtype=torch.float32)
        hist = torch.empty(S_max, B, T, device=self.device, dtype=torch.float32)
        steps = torch.zeros(1, device=self.device, dtype=torch.int32)
        self._check(self.L.taco_decode(self._h, mode, _ptr(memory), _ptr(lens), B, T, _ptr(tgt), F_tgt, S_max,
                                       _ptr(masks), C.c_uint64(int(philox_seed) & (2 ** 64 - 1)), int(bool(reset_state)),
                                       int(bool(logits_only)), _ptr(frames), _ptr(stop), _ptr(hist), _ptr(steps),
                                       self._stream()), "taco_decode")
        out = dict(frames=frames, stop=stop, align_hist=hist, S_max=S_max, steps=steps)
        if sync:
            out["steps"] = int(steps.item())
            if out["steps"] < 0:
                raise _lib.TacoError("taco_decode: the persistent decoder kernel aborted (an exchange poll timed out)")
        return out

    def alignments(self, align_hist, S):
        """[S_max, B, T] step-major history -> [B, T, S] (Tacotron.py:312)."""
        _, B, T = align_hist.shape
        out = torch.empty(B, T, S, device=self.device, dtype=torch.float32)
        self._check(self.L.taco_transpose_alignments(self._h, _ptr(align_hist), S, B, T, _ptr(out), self._stream()),
                    "taco_transpose_alignments")
        return out

    def postnet(self, frames, F):
        """raw frames [B, F_stride, 80] -> (decoder_output [B,F,80] clipped, mel_outputs [B,F,80])."""
        B, F_stride, M = frames.shape
        dec = torch.empty(B, F, M, device=self.device, dtype=torch.float32)
        mel = torch.empty(B, F, M, device=self.device, dtype=torch.float32)
        self._check(self.L.taco_postnet(self._h, _ptr(frames), B, F, F_stride, _ptr(dec), _ptr(mel), self._stream()),
                    "taco_postnet")
        return dec, mel

    def decoder_state(self, B, T):
        """TacotronDecoderCellState fields kept in the handle (Decoder.py:419-438)."""
        z = lambda n: torch.empty(B, n, device=self.device, dtype=torch.float32)
        c1, h1, c2, h2, att, al = z(256), z(256), z(256), z(256), z(768), z(T)
        self._check(self.L.taco_get_decoder_state(self._h, B, T, _ptr(c1), _ptr(h1), _ptr(c2), _ptr(h2), _ptr(att), _ptr(al),
                                                  self._stream()), "taco_get_decoder_state")
        return dict(c1=c1, h1=h1, c2=c2, h2=h2, attention=att, alignments=al)

    # ------------------------------------------------------------------ whole path
    def forward(self, ids, lens, mel_targets=None, reference_mel=None, gst_weights=None, max_iters=None,
                prenet_masks=None, philox_seed=0):
        """Encoder -> decoder -> postnet for one tower (inference / GTA modes).

        Inputs may live on host or device; outputs are device tensors shaped like the reference's
        tower outputs: decoder_output / mel_outputs [B, 2S, 80], stop [B, 2S], alignments [B, T, S]."""
        B = ids.shape[0]
        if B > MAX_BATCH_PER_CALL:
            raise ValueError("forward() handles <= %d utterances per call; use sharding.synthesize_batches" % MAX_BATCH_PER_CALL)
        r = self.hp.outputs_per_step
        memory = self.encode(ids, lens, reference_mel, gst_weights)
        d = self.decode(memory, lens, mel_targets, max_iters, prenet_masks, philox_seed)
        S = d["steps"]
        dec, mel = self.postnet(d["frames"], S * r)
        return dict(encoder_outputs=memory, decoder_output=dec, mel_outputs=mel, stop=d["stop"][:, : S * r],
                    alignments=self.alignments(d["align_hist"], S), steps=S, frames_raw=d["frames"][:, : S * r])

    # ------------------------------------------------------------------ host-buffer entry (the e2e path)
    def _pinned(self, key, shape, dtype):
        cache = self.__dict__.setdefault("_pin_cache", {})
        t = cache.get(key)
        n = int(np.prod(shape)) if len(shape) else 1
        if t is None or t.dtype != dtype or t.numel() < n:
            t = torch.empty(max(n, 1), dtype=dtype, pin_memory=True)
            cache[key] = t
        return t[:n].view(*shape)

    def _h2d(self, key, arr, dtype):
        a = np.ascontiguousarray(arr)
        p = self._pinned(key, a.shape, dtype)
        p.copy_(torch.from_numpy(a).to(dtype))
        return p.to(self.device, non_blocking=True), p.numel() * p.element_size()

    def forward_host(self, ids, lens, mel_targets=None, reference_mel=None, gst_weights=None, max_iters=None,
                     prenet_masks=None, philox_seed=0, fetch=("mel_outputs", "alignments", "stop")):
        """``forward`` on HOST numpy arrays: inputs staged through pinned memory and copied host->device,
        the tensors the reference's ``session.run`` fetches (Tacotron_synthesizer.py:110: mel_outputs,
        alignments, stop tokens) copied device->host into pinned buffers.  Returns numpy views + byte counts."""
        h2d = 0
        ids_d, n = self._h2d("ids", ids, torch.int32); h2d += n
        lens_d, n = self._h2d("lens", lens, torch.int32); h2d += n
        tgt_d = ref_d = gw_d = None
        if mel_targets is not None:
            tgt_d, n = self._h2d("tgt", mel_targets, torch.float32); h2d += n
        if reference_mel is not None:
            ref_d, n = self._h2d("ref", reference_mel, torch.float32); h2d += n
        elif gst_weights is not None:
            gw_d, n = self._h2d("gw", gst_weights, torch.float32); h2d += n
        if prenet_masks is not None:
            prenet_masks, n = self._h2d("masks", prenet_masks, torch.uint8); h2d += n
        out = self.forward(ids_d, lens_d, tgt_d, ref_d, gw_d, max_iters, prenet_masks, philox_seed)
        res = {"steps": out["steps"]}
        d2h = 4  # the int32 step count read back by forward()
        for k in fetch:
            t = out[k].contiguous()
            p = self._pinned("out_" + k, tuple(t.shape), t.dtype)
            p.copy_(t, non_blocking=True)
            d2h += t.numel() * t.element_size()
            res[k] = p
        torch.cuda.current_stream(self.device).synchronize()
        for k in fetch:
            res[k] = res[k].numpy()
        res["h2d_bytes"], res["d2h_bytes"] = h2d, d2h
        return res


class EnginePool:
    """``in_flight`` engines (own library handle, workspaces, pinned buffers and CUDA stream each) on ONE GPU, driven by
    one host thread each, so that several independent utterance batches are in flight at once.

    Why: the persistent decoder kernel of a batch occupies 7 clusters x 16 SMs for ~12 ms and is latency-bound; the
    encoder / postnet / copies of the neighbouring batches run on the remaining SMs and copy engines meanwhile (decoder
    kernels themselves serialise: two do not fit).  Batches stay independent -- same results as one at a time.  Weights
    are replicated per engine (82 MB each)."""

    def __init__(self, hp, weights=None, device=0, precision="fp32", in_flight=2):
        if in_flight < 1:
            raise ValueError("in_flight must be >= 1")
        self.device = device
        self.engines = [TacoEngine(hp, weights, device=device, precision=precision) for _ in range(in_flight)]
        if in_flight > 1:
            # the cluster decoder of one batch leaves 36 SMs free: a BiLSTM of <= 32 CTAs (instead of 128) runs on them, so the
            # whole encoder of the next batch overlaps the decoder instead of waiting for it
            for e in self.engines:
                e.set_encoder_ctas(32)
        self.streams = [torch.cuda.Stream(device=device) for _ in range(in_flight)]

    def __len__(self):
        return len(self.engines)

    def close(self):
        for e in self.engines:
            e.close()
        self.engines = []

    @property
    def launch_count(self):
        return sum(e.launch_count for e in self.engines)

    def map(self, fn, items, after=None):
        """``[fn(engine, item) for item in items]`` with up to ``in_flight`` items running concurrently; results keep the
        order of ``items``.  ``after``: optional CUDA event every stream waits for first.  Returns after all streams are
        drained; the first exception of a worker is re-raised."""
        import threading
        items = list(items)
        results = [None] * len(items)
        lock = threading.Lock()
        state = {"next": 0, "error": None}

        def worker(i):
            try:
                torch.cuda.set_device(self.device)
                with torch.cuda.stream(self.streams[i]):
                    if after is not None:
                        self.streams[i].wait_event(after)
                    while True:
                        with lock:
                            k = state["next"]
                            if k >= len(items) or state["error"] is not None:
                                break
                            state["next"] = k + 1
                        results[k] = fn(self.engines[i], items[k])
                    self.streams[i].synchronize()
            except BaseException as e:  # noqa: BLE001 - handed to the caller
                with lock:
                    if state["error"] is None:
                        state["error"] = e

        threads = [threading.Thread(target=worker, args=(i,)) for i in range(min(len(self.engines), max(1, len(items))))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if state["error"] is not None:
            raise state["error"]
        return results
