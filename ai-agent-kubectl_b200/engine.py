"""Generation engine behind ``B200ChatModel``: random-init weights of a named
architecture packed for the CUDA kernels, and ``generate`` over the C ABI
(``kx_engine_create`` / ``kx_engine_generate`` in include/kx.h).

torch is used for device memory only (weights live in torch tensors; the library
receives raw pointers).  There is no CPU path: creating an engine without a CUDA
device raises.
"""
from __future__ import annotations

import ctypes
import logging
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _lib, tokenizer, weights as weights_mod
from .chat_model import Engine, SlotEngine
from .model_config import ModelConfig, get_config

c_int32, c_float, c_void_p = ctypes.c_int32, ctypes.c_float, ctypes.c_void_p


class KxModelDesc(ctypes.Structure):
    _fields_ = [(n, c_int32) for n in ("hidden", "layers", "n_q", "n_kv", "head_dim", "inter", "vocab", "max_pos")] + \
               [("eps", c_float), ("max_batch", c_int32), ("max_seq", c_int32), ("prefill_chunk", c_int32),
                ("tp_size", c_int32), ("tp_rank", c_int32), ("vocab_offset", c_int32)]


class KxLayerWeights(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("ln1", "wqkv", "bqkv", "wo", "ln2", "wgu", "wd")]


class KxModelWeights(ctypes.Structure):
    _fields_ = [("embed", c_void_p), ("lm_head", c_void_p), ("norm", c_void_p), ("rope", c_void_p),
                ("layers", ctypes.POINTER(KxLayerWeights))]


class KxGenStats(ctypes.Structure):
    _fields_ = [("prefill_ms", c_float), ("decode_ms", c_float), ("decode_steps", c_int32), ("launches", c_int32),
                ("h2d_bytes", ctypes.c_int64), ("d2h_bytes", ctypes.c_int64)]


_lib.declare("kx_engine_create", ctypes.c_int, [ctypes.POINTER(KxModelDesc), ctypes.POINTER(KxModelWeights),
                                                 ctypes.POINTER(c_void_p)])
_lib.declare("kx_engine_set_collectives", ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p])
_lib.declare("kx_engine_p2p_export", ctypes.c_int, [c_void_p, c_void_p])
_lib.declare("kx_engine_p2p_import", ctypes.c_int, [c_void_p, c_void_p])
_lib.declare("kx_engine_p2p_launches", ctypes.c_int64, [c_void_p])
_lib.declare("kx_engine_attn_tc_launches", ctypes.c_int64, [c_void_p])
_lib.declare("kx_engine_trace", ctypes.c_int, [c_void_p, c_void_p, ctypes.c_int])
_lib.declare("kx_engine_trace_name", ctypes.c_char_p, [c_void_p, ctypes.c_int])
_lib.declare("kx_engine_set_prefix", ctypes.c_int, [c_void_p, c_void_p, c_int32, c_void_p])
_lib.declare("kx_engine_set_eos", ctypes.c_int, [c_void_p, c_int32])
_lib.declare("kx_engine_destroy", None, [c_void_p])
_lib.declare("kx_engine_device_bytes", ctypes.c_int64, [c_void_p])
_lib.declare("kx_engine_session_open", ctypes.c_int, [c_void_p])
_lib.declare("kx_engine_session_admit", ctypes.c_int, [c_void_p, c_int32, c_int32, c_void_p, c_int32, c_int32,
                                                        ctypes.POINTER(KxGenStats)])
_lib.declare("kx_engine_session_step", ctypes.c_int, [c_void_p, c_int32, c_int32])
_lib.declare("kx_engine_session_poll", ctypes.c_int, [c_void_p, c_void_p, c_void_p])
_lib.declare("kx_engine_session_read", ctypes.c_int, [c_void_p, c_int32, c_int32, c_void_p, c_int32])
_lib.declare("kx_engine_session_release", ctypes.c_int, [c_void_p, c_int32, c_int32])
_lib.declare("kx_engine_session_live", c_int32, [c_void_p])
_lib.declare("kx_engine_generate", ctypes.c_int, [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p,
                                                   ctypes.POINTER(KxGenStats), c_void_p])


def pack_weights(cfg: ModelConfig, w: Dict, device, tp_rank: int = 0, tp_size: int = 1) -> Dict:
    """Logical tensors (weights.make_weights) -> the layouts of include/kx.h on `device`.
    With tp_size > 1 only this rank's shard is packed: q/k/v heads, gate/up rows and the
    vocabulary are split by rows (column-parallel), wo / wd by input columns (row-parallel)."""
    import torch

    def rows(t, n_units, unit):           # contiguous block of n_units/tp units of `unit` rows
        per = n_units // tp_size
        return t[tp_rank * per * unit:(tp_rank + 1) * per * unit]

    hd = cfg.head_dim
    head = w["embed"] if cfg.tied else w["lm_head"]
    out = {"embed": w["embed"].to(device), "norm": w["norm"].to(device), "layers": []}
    if tp_size == 1:
        out["lm_head"] = out["embed"] if cfg.tied else head.to(device)
    else:
        out["lm_head"] = rows(head, cfg.vocab, 1).contiguous().to(device)
    for i in range(cfg.layers):
        wg, wu = rows(w[f"wg.{i}"], cfg.inter, 1), rows(w[f"wu.{i}"], cfg.inter, 1)
        wgu = torch.stack([wg, wu], dim=1).reshape(2 * wg.shape[0], cfg.hidden)   # rows g0,u0,g1,u1,...
        qcols = slice(tp_rank * cfg.q_dim // tp_size, (tp_rank + 1) * cfg.q_dim // tp_size)
        icols = slice(tp_rank * cfg.inter // tp_size, (tp_rank + 1) * cfg.inter // tp_size)
        layer = {
            "ln1": w[f"ln1.{i}"].to(device), "ln2": w[f"ln2.{i}"].to(device),
            "wqkv": torch.cat([rows(w[f"wq.{i}"], cfg.n_q, hd), rows(w[f"wk.{i}"], cfg.n_kv, hd),
                               rows(w[f"wv.{i}"], cfg.n_kv, hd)], dim=0).contiguous().to(device),
            "bqkv": (torch.cat([rows(w[f"bq.{i}"], cfg.n_q, hd), rows(w[f"bk.{i}"], cfg.n_kv, hd),
                                rows(w[f"bv.{i}"], cfg.n_kv, hd)]).float().to(device) if cfg.qkv_bias else None),
            "wo": w[f"wo.{i}"][:, qcols].contiguous().to(device), "wgu": wgu.contiguous().to(device),
            "wd": w[f"wd.{i}"][:, icols].contiguous().to(device),
        }
        out["layers"].append(layer)
    return out


class NcclComm:
    """A NCCL communicator created through the C API of the libnccl torch already loaded,
    so that the engine (C++) can enqueue collectives on its own stream.  The unique id is
    shared through torch.distributed (any backend)."""

    def __init__(self, rank: int, world: int, device: int, group=None):
        self.group = group
        import glob
        import os
        import torch
        import torch.distributed as dist
        cands = glob.glob(os.path.join(os.path.dirname(os.path.dirname(torch.__file__)), "nvidia", "nccl", "lib", "libnccl.so*"))
        self.lib = ctypes.CDLL(cands[0] if cands else "libnccl.so.2")
        uid = (ctypes.c_byte * 128)()
        if rank == 0:
            assert self.lib.ncclGetUniqueId(ctypes.byref(uid)) == 0
        box = [bytes(uid)]
        src = dist.get_global_rank(group, 0) if group is not None else 0     # `src` is a GLOBAL rank: the group's first member
        dist.broadcast_object_list(box, src=src, group=group)
        ctypes.memmove(uid, box[0], 128)
        self.comm = c_void_p()

        class _Uid(ctypes.Structure):
            _fields_ = [("b", ctypes.c_byte * 128)]
        u = _Uid()
        ctypes.memmove(ctypes.byref(u), uid, 128)
        self.lib.ncclCommInitRank.argtypes = [ctypes.POINTER(c_void_p), ctypes.c_int, _Uid, ctypes.c_int]
        with torch.cuda.device(device):
            rc = self.lib.ncclCommInitRank(ctypes.byref(self.comm), world, u, rank)
        if rc != 0:
            raise RuntimeError(f"ncclCommInitRank failed with {rc}")
        self.allreduce = ctypes.cast(self.lib.ncclAllReduce, c_void_p)
        self.allgather = ctypes.cast(self.lib.ncclAllGather, c_void_p)

    def close(self):
        if self.comm:
            self.lib.ncclCommDestroy.argtypes = [c_void_p]
            self.lib.ncclCommDestroy(self.comm)
            self.comm = None


class Generator:
    """One model replica on one GPU."""

    def __init__(self, arch: str, seed: int = 3, max_batch: int = 64, max_seq: int = 320, device: int = 0,
                 weights: Optional[Dict] = None, prefill_chunk: int = 32768, tp_rank: int = 0, tp_size: int = 1,
                 nccl: Optional["NcclComm"] = None, p2p: bool = True, embed_std: Optional[float] = None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("the generator needs a CUDA device; there is no CPU fallback")
        self.cfg = get_config(arch)
        self.device = torch.device("cuda", device)
        self.max_batch, self.max_seq = max_batch, max_seq
        if weights is not None:
            self.logical = weights
        elif embed_std is not None:      # per-tensor seeds: regenerate on demand, keep nothing on the host (8B: 16 GB otherwise)
            self.logical = weights_mod.LazyWeights(self.cfg, seed, embed_std)
        else:
            self.logical = weights_mod.make_weights(self.cfg, seed)
        with torch.cuda.device(self.device):
            self.packed = pack_weights(self.cfg, self.logical, self.device, tp_rank, tp_size)
            self.rope = weights_mod.rope_table(self.cfg, self.cfg.max_pos).to(self.device)
            c = self.cfg
            if tp_size > 1 and (c.n_q % tp_size or c.n_kv % tp_size or c.inter % tp_size or c.vocab % (8 * tp_size)):
                raise ValueError(f"{arch} cannot be split {tp_size} ways")
            self.tp_rank, self.tp_size = tp_rank, tp_size
            self.vocab_local = c.vocab // tp_size
            desc = KxModelDesc(c.hidden, c.layers, c.n_q // tp_size, c.n_kv // tp_size, c.head_dim, c.inter // tp_size,
                               self.vocab_local, c.max_pos, c.eps, max_batch, max_seq, prefill_chunk,
                               tp_size, tp_rank, tp_rank * self.vocab_local)
            arr = (KxLayerWeights * c.layers)()
            for i, l in enumerate(self.packed["layers"]):
                arr[i] = KxLayerWeights(l["ln1"].data_ptr(), l["wqkv"].data_ptr(),
                                        l["bqkv"].data_ptr() if l["bqkv"] is not None else None, l["wo"].data_ptr(),
                                        l["ln2"].data_ptr(), l["wgu"].data_ptr(), l["wd"].data_ptr())
            self._layer_array = arr
            mw = KxModelWeights(self.packed["embed"].data_ptr(), self.packed["lm_head"].data_ptr(),
                                self.packed["norm"].data_ptr(), self.rope.data_ptr(), arr)
            h = c_void_p()
            _lib.check(_lib.lib().kx_engine_create(ctypes.byref(desc), ctypes.byref(mw), ctypes.byref(h)))
            self._h = h
            if tp_size > 1:
                if nccl is None:
                    raise ValueError("tensor parallelism needs an NcclComm")
                self.nccl = nccl
                _lib.check(_lib.lib().kx_engine_set_collectives(h, nccl.comm, nccl.allreduce, nccl.allgather))
                if tp_size == 2 and p2p:
                    self._setup_p2p(h, tp_rank, nccl.group)
        self.last_stats: Optional[KxGenStats] = None

    @staticmethod
    def _setup_p2p(h, tp_rank, group):
        """Swap CUDA IPC handles with the peer rank so the decode GEMMs can reduce over NVLink themselves."""
        import torch.distributed as dist
        mine = ctypes.create_string_buffer(128)
        _lib.check(_lib.lib().kx_engine_p2p_export(h, mine))
        box = [None, None]
        dist.all_gather_object(box, bytes(mine.raw), group=group)
        peer = ctypes.create_string_buffer(box[1 - tp_rank], 128)
        _lib.check(_lib.lib().kx_engine_p2p_import(h, peer))

    def p2p_launches(self) -> int:
        return int(_lib.lib().kx_engine_p2p_launches(self._h))

    def attn_tc_launches(self) -> int:
        """Prefill attention launches that took the tcgen05 kernel so far."""
        return int(_lib.lib().kx_engine_attn_tc_launches(self._h))

    def trace(self, max_slots: int = 2048):
        """Diagnostic (KX_DEBUG=trace): [(label, start, deps_ready, main_done, end, exchanged, ...)] in ns for every
        instrumented launch of the last decode step, in launch order."""
        buf = np.zeros((max_slots, 8), np.uint64)
        n = _lib.lib().kx_engine_trace(self._h, buf.ctypes.data, max_slots)
        if n < 0:
            _lib.check(n)
        return [(_lib.lib().kx_engine_trace_name(self._h, i).decode(),) + tuple(int(x) for x in buf[i]) for i in range(n)]

    def device_bytes(self) -> int:
        return int(_lib.lib().kx_engine_device_bytes(self._h))

    def set_prefix(self, tokens) -> None:
        """Prefill a shared prefix once (length a multiple of 16); later prompts continue it."""
        import torch
        toks = np.ascontiguousarray(tokens, np.int32).reshape(-1)
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(_lib.lib().kx_engine_set_prefix(self._h, toks.ctypes.data if toks.size else None, toks.size, st))

    def set_eos(self, eos_id: Optional[int]) -> None:
        """Token id that ends a sequence (None: none).  Decode stops once every sequence of the batch has emitted it."""
        _lib.check(_lib.lib().kx_engine_set_eos(self._h, -1 if eos_id is None else int(eos_id)))

    def generate_ids(self, tokens: np.ndarray, n_new: int, want_logits: bool = False):
        """tokens int32 [B, S] -> (ids int32 [B, n_new], logits torch fp32 [n_new, B, V] or None)."""
        import torch
        tokens = np.ascontiguousarray(tokens, np.int32)
        B, S = tokens.shape
        out = np.empty((B, n_new), np.int32)
        stats = KxGenStats()
        with torch.cuda.device(self.device):
            logits = torch.empty((n_new, B, self.vocab_local), dtype=torch.float32, device=self.device) if want_logits else None
            st = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(_lib.lib().kx_engine_generate(self._h, tokens.ctypes.data, B, S, n_new, out.ctypes.data,
                                                     logits.data_ptr() if want_logits else None, ctypes.byref(stats), st))
        self.last_stats = stats
        return out, logits

    # ---- continuous batching: thin wrappers of kx_engine_session_* (include/kx.h), driven by one host thread ----
    def session_open(self) -> None:
        _lib.check(_lib.lib().kx_engine_session_open(self._h))

    def session_admit(self, slot0: int, tokens: np.ndarray, max_new: int) -> KxGenStats:
        """tokens int32 [n, S] -> slots slot0 .. slot0+n-1 hold the prompts' KV and their first token."""
        tokens = np.ascontiguousarray(tokens, np.int32)
        n, S = tokens.shape
        stats = KxGenStats()
        _lib.check(_lib.lib().kx_engine_session_admit(self._h, slot0, n, tokens.ctypes.data, S, max_new, ctypes.byref(stats)))
        return stats

    def session_step(self, n_steps: int, n_slots: int) -> None:
        _lib.check(_lib.lib().kx_engine_session_step(self._h, n_steps, n_slots))

    def session_poll(self):
        """(count int32 [max_batch] (-1: free slot), finished int32 [max_batch]) after the enqueued steps."""
        cnt, fin = np.empty(self.max_batch, np.int32), np.empty(self.max_batch, np.int32)
        _lib.check(_lib.lib().kx_engine_session_poll(self._h, cnt.ctypes.data, fin.ctypes.data))
        return cnt, fin

    def session_read(self, slot0: int, n_slots: int, n: int) -> np.ndarray:
        """int32 [n_slots, n]: the first n tokens of the outputs of slots slot0 .. slot0+n_slots-1."""
        out = np.empty((n_slots, n), np.int32)
        _lib.check(_lib.lib().kx_engine_session_read(self._h, slot0, n_slots, out.ctypes.data, n))
        return out

    def session_release(self, slot0: int, n: int = 1) -> None:
        _lib.check(_lib.lib().kx_engine_session_release(self._h, slot0, n))

    def session_live(self) -> int:
        return int(_lib.lib().kx_engine_session_live(self._h))

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().kx_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


class PromptTooLong(RuntimeError):
    """The prompt does not fit the engine's context.  A RuntimeError on purpose: the reference maps ValueError to 422
    ("unsafe command", app.py:192-194) and everything else to 500 (app.py:195-197), which is where a context-length
    error of the OpenAI client ends up too."""


class B200Engine(SlotEngine):
    """``chat_model.Engine`` over a ``Generator``: prompts are byte-tokenised into ``prompt_len`` token slots (left-padded
    with spaces), ``max_new`` tokens are decoded greedily -- stopping early once every sequence has produced the
    end-of-sequence token -- and the ids are mapped back to text.

    A prompt longer than ``prompt_len`` tokens is an error for that request (``PromptTooLong``; the other requests of
    the batch are served).  ``on_overflow="truncate"`` keeps the prompt's tail instead: BASELINE configs 3-4 define their
    synthetic prompts as "padded / truncated to exactly 256 tokens", so bench.py asks for that explicitly.

    ``cancelled`` (one callable per prompt, from the batching front end): requests whose waiter is already gone when the
    batch is formed are not prefilled at all.  ``generate`` runs a static batch: a request cancelled after that point
    occupies its row until the batch ends.

    ``continuous=True`` makes ``B200ChatModel`` / ``EngineServer`` drive the engine through its slots instead
    (``chat_model.ContinuousFrontend`` over ``kx_engine_session_*``): requests join at the next step boundary, leave as
    soon as their own sequence ends, and a cancelled request frees its slot mid-sequence.  Single-GPU replicas only."""

    def __init__(self, arch: str, seed: int = 3, max_batch: int = 64, prompt_len: int = 256, max_new: int = 32,
                 device: int = 0, tp_rank: int = 0, tp_size: int = 1, nccl: Optional[NcclComm] = None,
                 prefix: Optional[str] = None, prefix_len: int = 0, auto_prefix: bool = False, max_prefix: int = 512,
                 parity_init: bool = False, p2p: bool = True, on_overflow: str = "error", eos: bool = True,
                 continuous: bool = False):
        if on_overflow not in ("error", "truncate"):
            raise ValueError("on_overflow must be 'error' or 'truncate'")
        if continuous and tp_size > 1:
            raise ValueError("continuous batching runs on single-GPU replicas (tp_size = 1)")
        self.continuous = continuous
        self.on_overflow = on_overflow
        self.auto_prefix, self.max_prefix = auto_prefix, (max_prefix // 16) * 16
        self._prefix_bytes, self.prefix_sets = b"", 0
        if auto_prefix:
            prefix_len = self.max_prefix
        self.gen = Generator(arch, seed=seed, max_batch=max_batch, max_seq=prefix_len + prompt_len + max_new, device=device,
                             tp_rank=tp_rank, tp_size=tp_size, nccl=nccl, p2p=p2p,
                             prefill_chunk=max(32768, prefix_len),
                             embed_std=get_config(arch).parity_embed_std if parity_init else None)
        if eos:
            self.gen.set_eos(tokenizer.EOS)
        self.max_batch, self.prompt_len, self.max_new = max_batch, prompt_len, max_new
        self.skipped_cancelled = 0
        if prefix_len and not auto_prefix:
            # constant instruction prefix (app.py:50-57 in the service; a system prompt in BASELINE config 5),
            # byte-tokenised and padded/truncated to a whole number of KV pages, prefilled once
            self.gen.set_prefix(np.asarray(tokenizer.encode_fixed(prefix or "", prefix_len), np.int32))

    def generate(self, prompts: Sequence[str], cancelled=None) -> List:
        """Completions in the order of ``prompts``; an element is an exception instance when that request alone failed
        (prompt too long) and ``""`` when its waiter had gone before the batch was formed."""
        out: List = [None] * len(prompts)
        live = []
        for i, p in enumerate(prompts):
            if cancelled is not None and cancelled[i] is not None and cancelled[i]():
                out[i] = ""
                self.skipped_cancelled += 1
            else:
                live.append(i)
        if self.auto_prefix and live:
            # requests that continue the cached prefix keep it; the others form a second batch that learns its own
            want = self._prefix_bytes
            first = [i for i in live if want and prompts[i].encode("utf-8").startswith(want)]
            taken = set(first)
            rest = [i for i in live if i not in taken]
            for part in (first, rest):
                if not part:
                    continue
                toks, keep = self._with_learned_prefix([prompts[i] for i in part], part, out)
                if keep:
                    ids, _ = self.gen.generate_ids(toks, self.max_new)
                    for i, text in zip(keep, tokenizer.decode_batch(ids)):
                        out[i] = text
            return out
        if live:
            room = self.prompt_len - 1                                   # one slot is the BOS token
            if self.on_overflow == "error":
                keep = []
                for i in live:
                    n = len(prompts[i].encode("utf-8"))
                    if n > room:
                        out[i] = PromptTooLong(f"prompt of {n + 1} tokens exceeds the engine's {self.prompt_len}-token prompt window")
                    else:
                        keep.append(i)
                live = keep
            toks = tokenizer.encode_fixed_batch([prompts[i] for i in live], self.prompt_len) if live else None
        if live:
            ids, _ = self.gen.generate_ids(toks, self.max_new)
            for i, text in zip(live, tokenizer.decode_batch(ids)):
                out[i] = text
        return out

    def _with_learned_prefix(self, prompts: Sequence[str], index: List[int], out: List):
        """Prefix cache for the service's own template: every prompt the chain renders starts with the
        same instruction text (app.py:50-57).  The longest common byte prefix of the batch (and of the
        prefix already cached) is rounded down to whole KV pages, prefilled once, and only the remainder
        of each prompt is prefilled per request.  A prompt whose remainder does not fit the suffix window is
        that request's error (or truncated, see ``on_overflow``); it does not shrink the cached prefix."""
        import os.path
        raw = [p.encode("utf-8") for p in prompts]
        common = os.path.commonprefix(raw + ([self._prefix_bytes] if self._prefix_bytes else []))
        page = 16
        n_tok = min(((1 + len(common)) // page) * page, self.max_prefix)      # BOS + bytes, whole pages
        want = common[:max(0, n_tok - 1)]
        if n_tok and want != self._prefix_bytes:
            self.gen.set_prefix(np.asarray([tokenizer.BOS] + list(want), np.int32))
            self._prefix_bytes, self.prefix_sets = want, self.prefix_sets + 1
        elif not n_tok and self._prefix_bytes:
            self.gen.set_prefix(np.zeros(0, np.int32))
            self._prefix_bytes = b""
        room = self.prompt_len if n_tok else self.prompt_len - 1
        keep_pos, keep_idx = [], []
        for k, (r, i) in enumerate(zip(raw, index)):
            n = len(r) - len(want if n_tok else b"")
            if n > room and self.on_overflow == "error":
                out[i] = PromptTooLong(f"prompt needs {n} tokens after the cached prefix, the suffix window holds {room}")
            else:
                keep_pos.append(k)
                keep_idx.append(i)
        if not keep_idx:
            return None, []
        if not n_tok:
            return tokenizer.encode_fixed_batch([prompts[k] for k in keep_pos], self.prompt_len), keep_idx
        return tokenizer.encode_suffix_batch([raw[k][len(want):] for k in keep_pos], self.prompt_len), keep_idx

    # ---- SlotEngine (continuous batching), called from ContinuousFrontend's thread ----
    def slots_open(self) -> None:
        self.gen.session_open()

    def slots_prepare(self, prompts: Sequence[str], idle: bool) -> List:
        out: List = [None] * len(prompts)
        index = list(range(len(prompts)))
        if not self.auto_prefix:
            room = self.prompt_len - 1
            keep = []
            for i in index:
                n = len(prompts[i].encode("utf-8"))
                if n > room and self.on_overflow == "error":
                    out[i] = PromptTooLong(f"prompt of {n + 1} tokens exceeds the engine's {self.prompt_len}-token prompt window")
                else:
                    keep.append(i)
            toks = tokenizer.encode_fixed_batch([prompts[i] for i in keep], self.prompt_len) if keep else []
        elif idle and not (self._prefix_bytes and any(p.encode("utf-8").startswith(self._prefix_bytes) for p in prompts)):
            # nothing is decoding and nobody continues the cached prefix: it is learned again from these prompts
            # (which ends the session)
            sets = self.prefix_sets
            had = self._prefix_bytes
            toks, keep = self._with_learned_prefix(prompts, index, out)
            if self.prefix_sets != sets or self._prefix_bytes != had:
                self.gen.session_open()
            toks = toks if keep else []
        else:
            # sequences are decoding against the cached prefix (or some of these prompts continue it): only prompts that
            # continue it are admitted, the others wait until the engine is idle and nobody needs this prefix
            want = self._prefix_bytes
            raw = [p.encode("utf-8") for p in prompts]
            room = self.prompt_len if want else self.prompt_len - 1
            keep = []
            for i in index:
                if want and not raw[i].startswith(want):
                    continue                                             # None: wait until the engine is idle
                n = len(raw[i]) - len(want)
                if n > room and self.on_overflow == "error":
                    out[i] = PromptTooLong(f"prompt needs {n} tokens after the cached prefix, the suffix window holds {room}")
                else:
                    keep.append(i)
            if not keep:
                toks = []
            elif want:
                toks = tokenizer.encode_suffix_batch([raw[i][len(want):] for i in keep], self.prompt_len)
            else:
                toks = tokenizer.encode_fixed_batch([prompts[i] for i in keep], self.prompt_len)
        for i, row in zip(keep, toks):
            out[i] = row
        return out

    def slots_admit(self, slot0: int, rows) -> None:
        self.last_admit = self.gen.session_admit(slot0, np.stack(rows), self.max_new)

    def slots_step(self, n_steps: int, n_slots: int) -> None:
        self.gen.session_step(n_steps, n_slots)

    def slots_poll(self):
        return self.gen.session_poll()

    def slots_read(self, slot0: int, counts: Sequence[int]) -> List[str]:
        ids = self.gen.session_read(slot0, len(counts), max(1, max(counts)))
        for j, c in enumerate(counts):
            ids[j, c:] = tokenizer.EOS                                  # beyond what the slot produced
        return tokenizer.decode_batch(ids)

    def slots_release(self, slot0: int, n: int) -> None:
        self.gen.session_release(slot0, n)

    def close(self):
        self.gen.close()


def default_engine_factory(arch: str) -> Engine:
    """What ``B200ChatModel(...)`` starts when the reference constructs its ``llm`` (app.py:108-117).
    The service's prompts share the template's instruction text: its KV is cached, only the rest is prefilled.
    ``KX_DEVICE`` picks the GPU (default 0); ``KX_REPLICAS=n`` starts n engines on GPUs KX_DEVICE .. KX_DEVICE+n-1
    inside this process behind a ReplicaRouter (hash affinity on the cache key), so one unmodified ``app.py``
    worker drives the whole box; alternatively run one uvicorn worker per GPU with its own KX_DEVICE.
    ``KX_ENGINE_SOCKET=/path``: the engine lives in another process (``python -m kubectl_b200.engine_server``) shared
    by all uvicorn workers of the service; this process gets a forwarding ``RemoteEngine``."""
    import os
    sock = os.environ.get("KX_ENGINE_SOCKET")
    if sock:   # several uvicorn workers in front of one engine process (engine_server.py): this worker only forwards
        from .engine_server import RemoteEngine
        return RemoteEngine(sock)
    dev0, n = int(os.environ.get("KX_DEVICE", "0")), int(os.environ.get("KX_REPLICAS", "1"))

    logging.getLogger(__name__).warning(
        "B200ChatModel serves seeded RANDOM-INIT weights of the %s architecture through a byte-level tokenizer (no checkpoint or "
        "tokenizer vocabulary can be loaded offline): completions are noise and the reference's validator will reject them (422).",
        arch)

    def make(device: int, continuous: bool) -> Engine:
        return B200Engine(arch, prompt_len=128, auto_prefix=True, device=device, continuous=continuous)
    if n <= 1:
        return make(dev0, True)        # one replica: continuous batching (requests join and leave at step boundaries)
    from .router import ReplicaRouter
    return ReplicaRouter([make(dev0 + i, False) for i in range(n)])   # the router hands each replica static batches
