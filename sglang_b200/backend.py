"""B200AttnBackend: the `--attention-backend b200` plugin.

Implements the reference AttentionBackend API (srt/layers/attention/base_attn_backend.py:21-281)
over the unchanged ForwardBatch, ReqToTokenPool page table and MHATokenToKVPool:

    init_forward_metadata{,_out_graph,_in_graph}   one planner launch per forward, on the device
    forward_decode                                 KV-store scatter + split-KV decode + merge
    forward_extend                                 KV-store scatter + prefix/causal extend
    init_cuda_graph_state / get_cuda_graph_seq_len_fill_value   CUDA-graph contract

Behavioural model: torch_native_backend.py:279-402 (what is computed) and
triton_backend.py:572-634, 962-1063, 1250-1437, 1714-1864 (metadata + graph hooks).
All device work is hand-written sm_100a CUDA behind include/b200attn.h; there is no Triton,
no torch fallback, and construction fails on a non-CUDA device.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops

try:  # inside an SGLang process: subclass the real ABC so isinstance checks hold
    from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as _Base  # type: ignore
except Exception:  # standalone (tests, bench, GPU box): the mirror of the same API
    from .srt_mirror.base_attn_backend import AttentionBackend as _Base


def _attn_tp_size(model_runner) -> int:
    try:
        from sglang.srt.runtime_context import get_parallel  # type: ignore

        return int(get_parallel().attn_tp_size)
    except Exception:
        return int(getattr(model_runner, "attn_tp_size", 1) or 1)


def _spec_num_draft_tokens(model_runner) -> int:
    """`--speculative-num-draft-tokens` (triton_backend.py:171), 0 without speculative decoding."""
    try:
        from sglang.srt.runtime_context import get_spec  # type: ignore

        n = get_spec().speculative_num_draft_tokens
    except Exception:
        n = getattr(getattr(model_runner, "server_args", None), "speculative_num_draft_tokens", None)
    return int(n or 0)


class _ExtendState:
    """One set of extend-side device buffers: planner work list, per-request length vectors for the
    speculative modes (int32 and int64 flavours, matching `seq_lens`), custom-mask bytes and their indptr."""

    def __init__(self, ws, max_bs, mask, device):
        self.ws, self.max_bs, self.max_tiles, self.mask = ws, max_bs, ws.max_tiles, mask
        self.device = device
        # split-KV partials of one-tile plans: grown on demand; a buffer a captured graph may hold is never freed
        self.split_ws = None
        self._split_keep = []
        self._ext = {dt: torch.zeros(max_bs, dtype=dt, device=device) for dt in (torch.int32, torch.int64)}
        self._pre = {dt: torch.zeros(max_bs, dtype=dt, device=device) for dt in (torch.int32, torch.int64)}
        self.mask_indptr = torch.zeros(max_bs + 1, dtype=torch.int64, device=device)

    @staticmethod
    def create(max_bs, max_tiles, mask_bytes, device, num_ctas) -> "_ExtendState":
        ws = ops.ExtendWorkspace.create(max_bs, max_tiles, device, num_ctas=num_ctas)
        mask = torch.zeros(max(int(mask_bytes), 1), dtype=torch.uint8, device=device)
        return _ExtendState(ws, max_bs, mask, device)

    def fits(self, bs, tiles, mask_bytes) -> bool:
        return bs <= self.max_bs and tiles <= self.max_tiles and mask_bytes <= self.mask.numel()

    def split_buffer(self, floats: int) -> torch.Tensor:
        if self.split_ws is None or self.split_ws.numel() < floats:
            self.split_ws = torch.empty(int(floats), dtype=torch.float32, device=self.device)
            self._split_keep.append(self.split_ws)
        return self.split_ws

    def ext_lens(self, dtype):
        return self._ext[dtype]

    def pre_lens(self, dtype):
        return self._pre[dtype]


class B200AttnBackend(_Base):
    #: rows of the fast decode kernel's work list per persistent CTA
    DECODE_WAVES = 16

    def __init__(self, model_runner, skip_prefill: bool = False):
        super().__init__()
        dev = torch.device(model_runner.device)
        if dev.type != "cuda":
            raise RuntimeError("B200AttnBackend needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = dev
        sm_count, _, _ = ops.device_info(dev)
        self.sm_count = sm_count
        # pools are resolved through the backend by other code (forward_context.py:70-76)
        self.req_to_token_pool = model_runner.req_to_token_pool
        self.token_to_kv_pool = model_runner.token_to_kv_pool
        self.page_size = int(getattr(model_runner, "page_size", 1) or 1)
        self.skip_prefill = skip_prefill

        cfg = model_runner.model_config
        tp = _attn_tp_size(model_runner)
        if hasattr(cfg, "get_max_num_attention_heads"):
            self.num_head = cfg.get_max_num_attention_heads() // tp
        else:
            self.num_head = cfg.num_attention_heads // tp
        self.num_kv_head = cfg.get_num_kv_heads(tp)
        pool = self.token_to_kv_pool
        if not (isinstance(getattr(pool, "k_buffer", None), list) and pool.k_buffer[0].dim() == 3):
            raise NotImplementedError(
                "b200 backend supports the NHD MHATokenToKVPool layout [slots, Hkv, D] only")
        self.head_dim = pool.k_buffer[0].shape[-1]
        self.v_head_dim = pool.v_buffer[0].shape[-1]
        self.kv_dtype = pool.dtype
        self._q_dtype = getattr(model_runner, "dtype", torch.bfloat16)
        # Sliding-window layers (hybrid SWA models): layer.sliding_window_size = W means the query at position x
        # attends positions [x - W, x] -- W + 1 keys, the rule of torch_native_backend.py:36-48 and of the Triton
        # extend kernel (extend_attention.py:391-398).  The Triton *decode* metadata clamps to min(seq_len, W) keys
        # instead (triton_backend.py:686-691); this backend follows the torch_native oracle in both modes.
        sw = getattr(model_runner, "sliding_window_size", None)
        self.sliding_window_size = int(sw) if sw is not None and sw > -1 else None
        if hasattr(pool, "translate_loc_from_full_to_swa"):
            raise NotImplementedError(
                "b200 backend: the separate SWA KV pool (SWAKVPool) is not supported -- launch hybrid sliding-window "
                "models with --disable-hybrid-swa-memory so that every layer uses the full MHATokenToKVPool "
                "(sliding-window layers then read their window from it)")

        self.max_bs = int(self.req_to_token_pool.size)
        self.max_context_len = int(getattr(cfg, "context_len", self.req_to_token_pool.max_context_len))
        self.decode_kernel = ops.KERNEL_AUTO
        self.extend_kernel = ops.KERNEL_AUTO
        self.decode_ctas = sm_count * ops.decode_ctas_per_sm(self.head_dim, self.kv_dtype)
        self._decode_items = ops.decode_items_per_chunk(self.num_head, self.num_kv_head, self.head_dim,
                                                        self.v_head_dim, self._q_dtype, self.kv_dtype)
        self.decode_ws = ops.DecodeWorkspace.create(
            self.max_bs, self.num_head, self.v_head_dim, dev, num_ctas=self.decode_ctas, waves=self.DECODE_WAVES)
        # second split-KV plan for the sliding-window layers (same partial buffers: layers run one after another)
        self.decode_ws_swa = self._swa_workspace(self.decode_ws)
        # Extend-side state comes in two sets (see _ExtendState): `_ext_graph` is allocated once by
        # init_cuda_graph_state and never touched again (its addresses are baked into captured graphs);
        # `_ext_eager` grows on demand for eager batches that do not fit it.  `_ext` is the set the current
        # forward uses, chosen in init_forward_metadata_out_graph.
        self._ext_graph: Optional[_ExtendState] = None
        self._ext_eager: Optional[_ExtendState] = None
        self._ext: Optional[_ExtendState] = None
        self._ext_captured = {}        # (mode, padded bs) -> the set a captured graph holds
        self._ext_lens = None          # (prefix_lens, extend_lens, custom_mask, mask_indptr) of this forward
        self._graph_state_frozen = False
        group = self.num_head // self.num_kv_head
        self._extend_block_q = ops.extend_block_q(self.head_dim, self.v_head_dim, group, self._q_dtype,
                                                  self.kv_dtype, self.extend_kernel)
        # short extends (TARGET_VERIFY, DRAFT_EXTEND, a few new tokens behind a cached prefix): when no request
        # extends by more than one 128-row tile of tokens, plan one-tile items -- the tcgen05 kernel then runs its
        # one-tile instance instead of carrying an empty second tile (0 = that instance does not apply)
        self._one_tile_block_q = ops.extend_block_q_one_tile(self.head_dim, self.v_head_dim, group, self._q_dtype,
                                                             self.kv_dtype, self.extend_kernel)
        self._ext_block_q = self._extend_block_q      # block_q of the current forward's plan
        self._ext_splits, self._ext_split_ws = 1, None      # split-KV choice of the current forward (one-tile plans)
        self._ext_tiles = 0                                   # ... and the exact tile count handed to the library
        # > 0: this instance is step (offset - 1) of a multi-step draft decode (see B200MultiStepDraftBackend)
        self.draft_step_offset = 0
        self._draft_seq_lens = {}
        self._draft_tree = None                               # tree drafting: the wrapper's per-branch page table
        self._ext_short = False                               # the host knows that no request extends by more than one tile
        self.num_draft_tokens = _spec_num_draft_tokens(model_runner)
        self.forward_metadata = None
        #: optional (start, stop) torch.cuda.Event pair re-recorded around the next decode attention kernel
        self.attn_events = None

    def _swa_workspace(self, ws):
        if self.sliding_window_size is None:
            return None
        return ops.DecodeWorkspace(max_bs=ws.max_bs, max_chunks=ws.max_chunks, num_ctas=ws.num_ctas,
                                   plan=torch.zeros_like(ws.plan), ws_o=ws.ws_o, ws_lse=ws.ws_lse)

    def _window_keys(self, layer) -> int:
        sw = getattr(layer, "sliding_window_size", -1)
        return int(sw) + 1 if sw is not None and sw > -1 else 0

    @staticmethod
    def _draft_token_num(forward_batch) -> int:
        spec = getattr(forward_batch, "spec_info", None)
        n = getattr(spec, "draft_token_num", None)
        if n is None:
            raise ValueError("TARGET_VERIFY needs forward_batch.spec_info.draft_token_num")
        return int(n)

    # ------------------------------------------------------------------ metadata
    @property
    def extend_ws(self):
        return self._ext.ws if self._ext is not None else None

    def _select_ext(self, key, bs: int, tokens: int, mask_bytes: int, in_capture: bool) -> "_ExtendState":
        """Pick the extend-side buffer set for this forward.

        A capture records which set it baked into the graph under `key` = (mode, padded batch size); every later
        call with that key (the replays, and eager batches of the same shape -- a forward never interleaves with
        a replay) gets the same set back, so the replay refills exactly the buffers the graph reads.  Sets that a
        graph holds are never reallocated or freed; other eager batches use the graph set when they
        fit it, else a separate eager set that grows geometrically."""
        tiles = tokens // self._extend_block_q + bs + 1
        st = self._ext_captured.get(key)
        if st is not None and st.fits(bs, tiles, mask_bytes):
            return st
        g = self._ext_graph
        if in_capture:
            if g is None or not g.fits(bs, tiles, mask_bytes):
                # init_cuda_graph_state under-sized it (e.g. no speculative config on the runner): allocate a
                # bigger set for this and later captures; older graphs keep theirs alive through _ext_captured
                dims = (g.max_bs, g.max_tiles, g.mask.numel()) if g else (0, 0, 0)
                g = _ExtendState.create(max(bs, dims[0]), max(tiles, dims[1]), max(mask_bytes, dims[2]),
                                        self.device, self.sm_count)
                self._ext_graph = g
            self._ext_captured[key] = g
            return g
        if g is not None and g.fits(bs, tiles, mask_bytes):
            return g
        e = self._ext_eager
        if e is None or not e.fits(bs, tiles, mask_bytes):
            old_bs, old_tiles, old_mask = (e.max_bs, e.max_tiles, e.mask.numel()) if e else (0, 0, 0)
            e = _ExtendState.create(
                max(self.max_bs, bs, old_bs),
                tiles if tiles <= old_tiles else max(tiles, 2 * old_tiles),
                mask_bytes if mask_bytes <= old_mask else max(mask_bytes, 2 * old_mask),
                self.device, self.sm_count)
            self._ext_eager = e
        return e

    def _spec_tokens_per_req(self, forward_batch) -> int:
        """Fixed query rows per request of the speculative graph modes (triton_backend.py:470-475,538-546)."""
        mode, spec = forward_batch.forward_mode, getattr(forward_batch, "spec_info", None)
        if mode.is_target_verify():
            return self._draft_token_num(forward_batch)
        n = getattr(spec, "num_tokens_per_req", None)
        if n is None or int(n) <= 0:
            n = self.num_draft_tokens
        if int(n) <= 0:
            raise ValueError("draft-extend under CUDA graphs needs spec_info.num_tokens_per_req or "
                             "server_args.speculative_num_draft_tokens")
        return int(n)

    def init_forward_metadata_out_graph(self, forward_batch, in_capture: bool = False):
        """Per-forward host-side prep, outside graph capture (base_attn_backend.py:59-79).

        Decode needs nothing: the planner kernel reads `seq_lens` on the device, so a replay only has to refresh
        the runner's static `seq_lens` / `req_pool_indices`.  Extend modes choose the buffer set and fill the
        per-request length / mask buffers the recorded launches read (the reference's Triton backend does the
        same refill in `_apply_cuda_graph_metadata`, triton_backend.py:572-634)."""
        mode = forward_batch.forward_mode
        if in_capture:
            self._graph_state_frozen = True
        if mode.is_decode_or_idle():
            rows = forward_batch.batch_size * (self._draft_tree.topk if self._draft_tree is not None else 1)
            if rows > self.decode_ws.max_bs:
                raise ValueError(f"batch size {forward_batch.batch_size} exceeds the decode workspace "
                                 f"({self.decode_ws.max_bs} requests)")
            return
        bs = int(forward_batch.batch_size)
        spec = getattr(forward_batch, "spec_info", None)
        eager_lens = getattr(forward_batch, "extend_prefix_lens", None) is not None and not in_capture
        if mode.is_target_verify():
            # every request extends by draft_token_num tokens over `seq_lens` cached ones (triton_backend.py:807-866)
            n = self._spec_tokens_per_req(forward_batch)
            self._ext_block_q = self._pick_block_q(n)
            mask = getattr(spec, "custom_mask", None)
            st = self._select_ext((int(mode), bs), bs, bs * n, 0 if mask is None else int(mask.numel()), in_capture)
            pre = forward_batch.seq_lens[:bs]
            ext = st.ext_lens(pre.dtype)[:bs]
            ext.fill_(n)
            mask_indptr = None
            if mask is not None:
                # the mask tensor of a replay is not the capture-time one: copy it into the persistent buffer
                # (triton_backend.py:507-514)
                m = mask.view(torch.uint8) if mask.dtype == torch.bool else mask
                st.mask[: m.numel()].copy_(m, non_blocking=True)
                mask_indptr = st.mask_indptr[: bs + 1]
                torch.cumsum((pre.to(torch.int64) + n) * n, dim=0, out=mask_indptr[1:])
                mask = st.mask
            self._ext, self._ext_lens = st, (pre, ext, mask, mask_indptr)
            self._plan_splits(bs, bs * n, self._kv_hint(forward_batch, bs, in_capture))
        elif not eager_lens and mode.is_draft_extend_v2():
            # CUDA-graph draft-extend: a fixed number of query rows per request; the cached prefix is
            # seq_lens - accepted tokens, clamped for padded rows (triton_backend.py:529-565)
            n = self._spec_tokens_per_req(forward_batch)
            self._ext_block_q = self._pick_block_q(n)
            st = self._select_ext((int(mode), bs), bs, bs * n, 0, in_capture)
            seq = forward_batch.seq_lens[:bs]
            ext = st.ext_lens(seq.dtype)[:bs]
            ext.fill_(n)
            pre = st.pre_lens(seq.dtype)[:bs]
            acc = getattr(spec, "extend_seq_lens_tensor", None)
            if acc is None:
                acc = getattr(spec, "num_accept_tokens", None)
            if acc is not None:
                torch.clamp(seq - acc[:bs].to(seq.dtype), min=0, out=pre)
            else:
                pre.copy_(seq)
            self._ext, self._ext_lens = st, (pre, ext, None, None)
            self._plan_splits(bs, bs * n, self._kv_hint(forward_batch, bs, in_capture))
        else:
            T = getattr(forward_batch, "extend_num_tokens", None)
            if T is None:
                T = forward_batch.out_cache_loc.shape[0]
            lens_cpu = getattr(forward_batch, "extend_seq_lens_cpu", None)      # host copy the scheduler keeps
            self._ext_block_q = self._pick_block_q(max(lens_cpu) if lens_cpu is not None and len(lens_cpu) else None)
            self._ext = self._select_ext((int(mode), bs), bs, int(T), 0, in_capture)
            self._ext_lens = (forward_batch.extend_prefix_lens, forward_batch.extend_seq_lens, None, None)
            n_ext = sum(1 for e in lens_cpu if int(e) > 0) if lens_cpu is not None and len(lens_cpu) == bs else 0
            self._plan_splits(bs, int(T), self._kv_hint(forward_batch, bs, in_capture), tiles=n_ext)

    @staticmethod
    def _kv_hint(forward_batch, bs: int, in_capture: bool) -> int:
        """Mean context length per request from the host-side `seq_lens_sum` (0 = unknown: capture batches hold fill
        values)."""
        total = getattr(forward_batch, "seq_lens_sum", None)
        if in_capture or bs <= 0 or not isinstance(total, int) or total <= 0:
            return 0
        return total // bs

    def _plan_splits(self, bs: int, tokens: int, kv_hint: int = 0, tiles: int = 0):
        """Split-KV for one-tile plans (speculative verify / draft extend / a few new tokens): bs x Hkv work items
        rarely fill the SMs, so every item's kv range is cut into parts (ops.extend_pick_splits; the device trims the
        count for short contexts).  Under capture it is chosen from the batch shape alone (a graph keeps its choice);
        eager batches add the mean context length (`seq_lens_sum`, a host int) as a hint."""
        self._ext_splits, self._ext_split_ws, self._ext_tiles = 1, None, 0
        if not self._ext_short or bs <= 0 or tokens <= 0:       # (D = 256 plans are always one-tile: long prefills stay unsplit)
            return
        tiles = int(tiles) if tiles > 0 else bs          # short plans: one tile per request that extends at all
        s = ops.extend_pick_splits(tiles * self.num_kv_head, self.sm_count, kv_hint)
        if s > 1:
            self._ext_splits, self._ext_tiles = s, tiles
            self._ext_split_ws = self._ext.split_buffer(
                ops.extend_split_ws_floats(s, tokens, self.num_head, self.v_head_dim))

    def init_forward_metadata_in_graph(self, forward_batch):
        """One planner launch (static shape, graph-recordable, no host sync, reads persistent buffers only)."""
        mode = forward_batch.forward_mode
        if mode.is_idle():
            return
        if mode.is_decode():
            rows = forward_batch.batch_size * (self._draft_tree.topk if self._draft_tree is not None else 1)
            _, _, seq_lens = self._decode_inputs(forward_batch, rows, refresh=True)
            ops.decode_plan(self.decode_ws, seq_lens, self.num_kv_head, waves=self.DECODE_WAVES,
                            items_per_chunk=self._decode_items)
            if self.decode_ws_swa is not None:
                ops.decode_plan(self.decode_ws_swa, seq_lens, self.num_kv_head, waves=self.DECODE_WAVES,
                                items_per_chunk=self._decode_items, window_keys=self.sliding_window_size + 1)
        else:
            pre, ext = self._ext_lens[0], self._ext_lens[1]
            # page_size >= 16: the plan also verifies that pages are runs of consecutive pool rows (the paged
            # allocator's layout), which lets forward_extend fetch K/V by TMA page boxes
            paged = self.page_size >= 16 and pre is not None and pre.dtype == ext.dtype
            ops.extend_plan(self._ext.ws, ext, self._ext_block_q,
                            page_size=self.page_size if paged else 1,
                            req_to_token=self.req_to_token_pool.req_to_token if paged else None,
                            req_pool_indices=forward_batch.req_pool_indices[: ext.shape[0]] if paged else None,
                            extend_prefix_lens=pre if paged else None)

    def _decode_inputs(self, forward_batch, rows: int, refresh: bool = False):
        """(page table, request rows, lengths) of a decode launch over `rows` query rows.  Tree drafting
        (B200MultiStepDraftBackend with topk > 1) substitutes the per-branch table its wrapper builds once per draft."""
        tree = self._draft_tree
        if tree is not None:
            return tree.table, tree.rpi[:rows], tree.seq[self.draft_step_offset - 1][:rows]
        return (self.req_to_token_pool.req_to_token, forward_batch.req_pool_indices[:rows],
                self._decode_seq_lens(forward_batch, rows, refresh))

    def _decode_seq_lens(self, forward_batch, bs: int, refresh: bool = False):
        """The lengths a decode step attends.  A plain backend reads `forward_batch.seq_lens`.  Step i of a multi-step
        EAGLE draft (B200MultiStepDraftBackend) attends the prefix plus the i + 1 tokens drafted so far, whose slots
        the worker has written behind the prefix in `req_to_token` (cache_locs.py:104-108 for topk 1), i.e.
        `seq_lens + offset`: kept in a persistent buffer that the graph-recordable metadata hook refreshes."""
        seq = forward_batch.seq_lens[:bs]
        if self.draft_step_offset <= 0:
            return seq
        buf = self._draft_seq_lens.get(seq.dtype)
        if buf is None or buf.numel() < bs:
            if self._graph_state_frozen and buf is not None:
                raise RuntimeError("b200 backend: draft-step length buffer would have to grow after a graph capture")
            buf = torch.zeros(max(bs, self.max_bs), dtype=seq.dtype, device=seq.device)
            self._draft_seq_lens[seq.dtype] = buf
        if refresh:
            torch.add(seq, self.draft_step_offset, out=buf[:bs])
        return buf[:bs]

    def _pick_block_q(self, max_extend_len) -> int:
        """Query tokens per work item for this forward: one-tile items when the host knows that no request extends
        by more than one tile's worth of tokens (every request then is exactly one item), else two-tile items."""
        self._ext_short = bool(self._one_tile_block_q and max_extend_len is not None
                               and 0 < int(max_extend_len) <= self._one_tile_block_q)
        return self._one_tile_block_q if self._ext_short else self._extend_block_q

    # ------------------------------------------------------------------ CUDA graphs
    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        """Allocate every buffer whose address a captured graph will hold (decode_cuda_graph_runner.py:318,444
        call this with the largest capture batch, usually far below the pool size).  Workspaces only ever GROW
        here, and nothing is reallocated once a capture has started."""
        max_bs, max_num_tokens = int(max_bs), int(max_num_tokens)
        if max_bs > self.decode_ws.max_bs:
            if self._graph_state_frozen:
                raise RuntimeError("b200 backend: init_cuda_graph_state asked for a larger batch after a graph "
                                   "was captured; captured graphs hold the current workspace addresses")
            self.max_bs = max_bs
            self.decode_ws = ops.DecodeWorkspace.create(
                max_bs, self.num_head, self.v_head_dim, self.device, num_ctas=self.decode_ctas,
                waves=self.DECODE_WAVES)
            self.decode_ws_swa = self._swa_workspace(self.decode_ws)
        g = self._ext_graph
        tiles = max_num_tokens // self._extend_block_q + max_bs + 1
        # custom (tree) masks of TARGET_VERIFY: draft tokens x (context + draft tokens) bytes per request, the
        # reference's bound (triton_backend.py:1013-1017)
        mask_bytes = max_num_tokens * self.max_context_len if self.num_draft_tokens > 0 else 0
        if g is None or not g.fits(max_bs, tiles, mask_bytes):
            self._ext_graph = _ExtendState.create(max_bs, tiles, mask_bytes, self.device, self.sm_count)

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    def update_verify_buffers_to_fill_after_draft(self, spec_info, cuda_graph_bs):
        """EAGLE with the overlapped plan stream prepares the verify metadata BEFORE the draft has produced the tree
        (eagle_worker_common.py:505-535): the mask that init_forward_metadata_out_graph copied was not the final one.
        The worker calls this hook after the draft; everything else this backend derives for TARGET_VERIFY depends
        on `seq_lens` and the draft-token count only, so the tree mask is the one buffer to refresh (the Triton
        backend lets the tree kernel write into its own `verify_mask` buffer and has nothing to redo,
        triton_backend.py:1208-1215; this backend keeps `verify_mask` = None, so the worker hands the mask over in
        `spec_info.custom_mask`)."""
        mask = getattr(spec_info, "custom_mask", None)
        lens = self._ext_lens
        if mask is None or lens is None or lens[2] is None:
            return
        m = mask.view(torch.uint8) if mask.dtype == torch.bool else mask
        if m.numel() > lens[2].numel():
            raise ValueError(f"b200 backend: tree mask of {m.numel()} bytes exceeds the verify mask buffer "
                             f"({lens[2].numel()} bytes) chosen by init_forward_metadata_out_graph")
        lens[2][: m.numel()].copy_(m, non_blocking=True)

    # ------------------------------------------------------------------ KV write
    def _save_kv(self, layer, cache_loc, k, v):
        """Scatter the new rows straight into the pool's layer buffers (same layout and
        cast rules as MHATokenToKVPool.set_kv_buffer, memory_pool.py:2305-2381; slot 0 is
        the sink for padded rows, kvcache.cuh:215)."""
        pool = self.token_to_kv_pool
        idx = layer.layer_id - getattr(pool, "start_layer", 0)
        same = k.dtype == pool.dtype
        if cache_loc.dim() != 1 or not cache_loc.is_contiguous():
            # the EAGLE draft worker hands step i a strided row of its [steps, bs] slot table
            # (eagle_worker_v2.py:563-632); set_kv_buffer takes any index tensor, the scatter kernel wants a dense one
            cache_loc = cache_loc.reshape(-1).contiguous()
        ops.store_kv(
            k, v, pool.k_buffer[idx], pool.v_buffer[idx], cache_loc,
            cache_dtype=pool.dtype,
            k_scale=None if same else layer.k_scale_float,
            v_scale=None if same else layer.v_scale_float,
            size_limit=pool.size + pool.page_size, skip_index=0,
        )

    def _check_layer(self, layer):
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 backend: cross attention is not supported")
        if int(getattr(layer, "xai_temperature_len", -1) or -1) > 0:
            raise NotImplementedError("b200 backend: xai_temperature_len is not supported")
        wk = self._window_keys(layer)
        if wk and (self.sliding_window_size is None or wk != self.sliding_window_size + 1):
            raise ValueError(f"b200 backend: layer window {wk - 1} differs from the model runner's "
                             f"sliding_window_size {self.sliding_window_size}")

    @staticmethod
    def _logit_cap(layer) -> float:
        # logit_capping_mod (reference srt/layers/attention/triton_backend.py:83-88): tanh is the only method
        method = getattr(layer, "logit_capping_method", "tanh") or "tanh"
        cap = float(getattr(layer, "logit_cap", 0.0) or 0.0)
        if cap > 0.0 and method != "tanh":
            raise NotImplementedError(f"b200 backend: logit capping method {method!r} is not supported")
        return cap

    @staticmethod
    def _sinks(kwargs, device):
        sinks = kwargs.get("sinks")
        if sinks is None:
            return None
        return sinks.to(device=device, dtype=torch.float32).contiguous()

    def _descale(self, layer):
        if getattr(layer, "k_scale", None) is not None and getattr(layer, "v_scale", None) is not None:
            return float(layer.k_scale_float), float(layer.v_scale_float)
        return 1.0, 1.0

    def _pool_buffers(self, layer):
        pool = self.token_to_kv_pool
        # fetched per call: the pool may re-create its buffers (memory saver / resume)
        return pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id)

    # ------------------------------------------------------------------ forward
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        self._check_layer(layer)
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        if layer.qk_head_dim != layer.v_head_dim:
            o = q.new_empty((q.shape[0], layer.tp_q_head_num * layer.v_head_dim))
        else:
            o = torch.empty_like(q)
        if save_kv_cache and k is not None and v is not None:
            self._save_kv(layer, forward_batch.out_cache_loc, k, v)
        k_buf, v_buf = self._pool_buffers(layer)
        k_descale, v_descale = self._descale(layer)
        bs = q.shape[0]
        window_keys = self._window_keys(layer)
        rtt, rpi, seq_lens = self._decode_inputs(forward_batch, bs)
        ops.decode_attention(
            q.view(bs, layer.tp_q_head_num, layer.qk_head_dim),
            o.view(bs, layer.tp_q_head_num, layer.v_head_dim),
            k_buf, v_buf,
            rtt, rpi, seq_lens,
            self.decode_ws_swa if window_keys else self.decode_ws, layer.scaling,
            kv_dtype=self.kv_dtype, k_scale=k_descale, v_scale=v_descale,
            logit_cap=self._logit_cap(layer), kernel=self.decode_kernel,
            attn_events=self.attn_events, window_keys=window_keys, sinks=self._sinks(kwargs, q.device),
        )
        return o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        """EXTEND / MIXED / DRAFT_EXTEND: prefix + causal new tokens.  TARGET_VERIFY: every request extends by
        `spec_info.draft_token_num` tokens over `seq_lens` cached ones under `spec_info.custom_mask` (the tree
        mask replaces the causal rule on the new tokens, triton_backend.py:807-866).  With
        `forward_batch.mha_return_lse` the call returns `(o [T, Hq, Dv], lse [T, Hq])`
        (radix_attention.py:244-277)."""
        self._check_layer(layer)
        attn_out = getattr(forward_batch, "_attn_output", None)
        if attn_out is not None:
            o = attn_out
        elif layer.qk_head_dim != layer.v_head_dim:
            o = q.new_empty((q.shape[0], layer.tp_q_head_num * layer.v_head_dim))
        else:
            o = torch.empty_like(q)
        if save_kv_cache and k is not None and v is not None:
            self._save_kv(layer, forward_batch.out_cache_loc, k, v)
        k_buf, v_buf = self._pool_buffers(layer)
        k_descale, v_descale = self._descale(layer)
        from .srt_mirror.radix_attention import AttentionType as _AT

        attn_type = getattr(layer, "attn_type", None)
        causal = not (getattr(attn_type, "value", attn_type) == _AT.ENCODER_ONLY.value)
        T = q.shape[0]
        # chosen by init_forward_metadata_out_graph: persistent buffers under CUDA graphs
        prefix_lens, extend_lens, custom_mask, mask_indptr = self._ext_lens
        bs = extend_lens.shape[0]
        lse = None
        if getattr(forward_batch, "mha_return_lse", False):
            lse = torch.empty((T, layer.tp_q_head_num), dtype=torch.float32, device=q.device)
        ops.extend_attention(
            q.view(T, layer.tp_q_head_num, layer.qk_head_dim),
            o.view(T, layer.tp_q_head_num, layer.v_head_dim),
            k_buf, v_buf,
            self.req_to_token_pool.req_to_token,
            forward_batch.req_pool_indices[:bs], prefix_lens, extend_lens,
            self._ext.ws, layer.scaling,
            kv_dtype=self.kv_dtype, k_scale=k_descale, v_scale=v_descale,
            logit_cap=self._logit_cap(layer), causal=causal,
            kernel=self.extend_kernel, lse=lse,
            window_keys=self._window_keys(layer) if causal else 0, sinks=self._sinks(kwargs, q.device),
            custom_mask=custom_mask, mask_indptr=mask_indptr, mask_skip_prefix=True, page_size=self.page_size,
            num_splits=self._ext_splits, split_ws=self._ext_split_ws, num_tiles_hint=self._ext_tiles,
        )
        if lse is not None:
            return o.view(T, layer.tp_q_head_num, layer.v_head_dim), lse
        return o

    def support_triton(self):
        return True


class _DraftTree:
    """Per-branch page table of an EAGLE tree draft (topk > 1) and the per-step lengths: built once per draft by
    `refresh` (graph-recordable), shared by the per-step backends."""

    def __init__(self, topk: int, num_steps: int, page_size: int, device):
        self.topk, self.num_steps, self.page_size, self.device = int(topk), int(num_steps), int(page_size), device
        self.table = self.rpi = self.seq = None
        self.frozen = False

    def ensure(self, rows: int, width: int):
        if self.table is not None and self.table.shape[0] >= rows and self.table.shape[1] >= width:
            return
        if self.frozen:
            raise RuntimeError("b200 backend: the draft table would have to grow after a CUDA-graph capture")
        self.table = torch.zeros(rows, width, dtype=torch.int32, device=self.device)
        self.rpi = torch.arange(rows, dtype=torch.int64, device=self.device)
        self.seq = torch.ones(max(self.num_steps - 1, 1), rows, dtype=torch.int32, device=self.device)

    def refresh(self, forward_batch, req_to_token):
        bs = int(forward_batch.batch_size)
        rows = bs * self.topk
        self.ensure(rows, req_to_token.shape[1])
        seq = forward_batch.seq_lens[:bs]
        ops.build_draft_table(self.table, req_to_token, forward_batch.req_pool_indices[:bs], seq, self.topk,
                              self.num_steps, self.page_size)
        base = seq.to(torch.int32).repeat_interleave(self.topk)
        for i in range(self.num_steps - 1):
            torch.add(base, i + 1, out=self.seq[i][:rows])


class B200MultiStepDraftBackend:
    """EAGLE draft decode over several consecutive steps with `--speculative-eagle-topk 1` (chain drafting): the
    reference wraps `speculative_num_steps - 1` per-step backends in one object that the draft worker initialises once
    per draft and indexes per step (`TritonMultiStepDraftBackend`, triton_backend.py:1867-2042;
    eagle_worker_v2.py:608-650 runs step i under `attn_backends[i]`).  The Triton version materialises per-step
    kv_indices (`generate_draft_decode_kv_indices`); for topk 1 those are the request's `req_to_token` row up to
    `seq_len + i + 1`, which this backend's decode kernels read directly -- so a step is a plain decode whose lengths
    are `seq_lens + i + 1`.  Tree drafting (topk > 1): branch t attends the prefix plus the slots drafted on its own
    branch, a non-contiguous slice of the row -- `_DraftTree` builds one page-table row per (request, branch) once per
    draft (`b200attn_draft_table`, the dynamic-length counterpart of the reference's per-step kv_indices copy), and the
    steps are plain decodes over that table with num_seqs x topk rows."""

    needs_cpu_seq_lens: bool = False

    def __init__(self, model_runner, topk: int, speculative_num_steps: int):
        self.topk = int(topk)
        if self.topk < 1:
            raise ValueError("B200MultiStepDraftBackend: topk must be >= 1")
        self.speculative_num_steps = int(speculative_num_steps)
        self.device = model_runner.device
        self.attn_backends = []
        page_size = int(getattr(model_runner, "page_size", 1) or 1)
        # chain drafting reads the request's own req_to_token row; tree drafting needs one row per branch
        self._tree = _DraftTree(self.topk, self.speculative_num_steps, page_size, self.device) if self.topk > 1 else None
        for i in range(self.speculative_num_steps - 1):
            b = B200AttnBackend(model_runner, skip_prefill=True)
            b.draft_step_offset = i + 1
            if self._tree is not None:                      # a launch has num_seqs x topk query rows
                b._draft_tree = self._tree
                b.max_bs = b.max_bs * self.topk
                b.decode_ws = ops.DecodeWorkspace.create(b.max_bs, b.num_head, b.v_head_dim, b.device,
                                                         num_ctas=b.decode_ctas, waves=b.DECODE_WAVES)
                b.decode_ws_swa = b._swa_workspace(b.decode_ws)
            self.attn_backends.append(b)
        self.max_context_len = self.attn_backends[0].max_context_len if self.attn_backends else None
        self.req_to_token_pool = model_runner.req_to_token_pool
        self.token_to_kv_pool = model_runner.token_to_kv_pool

    @staticmethod
    def _decode_view(forward_batch):
        """The fields the per-step backends read, with the mode pinned to DECODE (forward_batch_info.py:1618-1656)."""
        from types import SimpleNamespace

        from .srt_mirror.forward_batch_info import ForwardMode as _FM

        mode = forward_batch.forward_mode
        try:                                   # inside an SGLang process the reference's own enum
            from sglang.srt.model_executor.forward_batch_info import ForwardMode as _RealFM  # type: ignore

            decode = _RealFM.DECODE
        except Exception:
            decode = _FM.DECODE
        if mode.is_decode():
            return forward_batch
        return SimpleNamespace(
            batch_size=forward_batch.batch_size, forward_mode=decode, actual_forward_mode=mode,
            req_pool_indices=forward_batch.req_pool_indices, seq_lens=forward_batch.seq_lens,
            seq_lens_sum=getattr(forward_batch, "seq_lens_sum", None),
            seq_lens_cpu=getattr(forward_batch, "seq_lens_cpu", None),
            out_cache_loc=getattr(forward_batch, "out_cache_loc", None),
            positions=getattr(forward_batch, "positions", None), spec_info=getattr(forward_batch, "spec_info", None))

    def _refresh_tree(self, forward_batch):
        if self._tree is not None and self.attn_backends:
            self._tree.refresh(forward_batch, self.req_to_token_pool.req_to_token)

    def init_forward_metadata(self, forward_batch):
        view = self._decode_view(forward_batch)
        self._refresh_tree(view)
        for b in self.attn_backends:
            b.init_forward_metadata(view)

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        if self._tree is not None:
            self._tree.ensure(int(max_bs) * self.topk, self.req_to_token_pool.req_to_token.shape[1])
        for b in self.attn_backends:
            b.init_cuda_graph_state(int(max_bs) * self.topk, max_num_tokens)
            n = max(int(max_bs), b.max_bs)
            for dt in (torch.int32, torch.int64):       # the per-step length buffers a captured graph will hold
                if dt not in b._draft_seq_lens or b._draft_seq_lens[dt].numel() < n:
                    b._draft_seq_lens[dt] = torch.zeros(n, dtype=dt, device=self.device)

    def init_forward_metadata_out_graph(self, forward_batch, in_capture: bool = False):
        view = self._decode_view(forward_batch)
        if self._tree is not None:
            self._tree.ensure(int(view.batch_size) * self.topk, self.req_to_token_pool.req_to_token.shape[1])
            self._tree.frozen = self._tree.frozen or in_capture
        for b in self.attn_backends:
            b.init_forward_metadata_out_graph(view, in_capture=in_capture)

    def init_forward_metadata_in_graph(self, forward_batch):
        view = self._decode_view(forward_batch)
        self._refresh_tree(view)
        for b in self.attn_backends:
            b.init_forward_metadata_in_graph(view)

    def get_cuda_graph_seq_len_fill_value(self):
        return 1
