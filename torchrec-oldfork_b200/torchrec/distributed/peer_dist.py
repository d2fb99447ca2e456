"""Fused lookup + exchange over NVLink peer memory for ShardedEmbeddingBagCollection.

The reference runs lookup -> NCCL collective -> re-layout copy for every sharding type
(table-wise: alltoall_pooled, comm_ops.py:203-560; row-wise: reduce_scatter_pooled, :382-890) and the
transposed collectives in backward.  Here every rank maps its peers' output / gradient buffers
(torch symmetric memory: CUDA VMM handles exchanged once) and the lookup kernels do the exchange:

forward   table-wise : b200_tbe_forward_peer stores the pooled row of global sample b straight into
                       rank (b // B_local)'s final [B_local, sum D] buffer, at that feature's column.
          row-wise   : the same kernel stores the partial sums of its NON-EMPTY bags into slot `my rank`
                       of the destination's (all-zero) staging buffer; after the barrier
                       b200_peer_stage_reduce sums the W slots in rank order (deterministic) into the
                       final buffer and zeroes what it read.  With one id per bag this moves 1/W of
                       what a reduce-scatter of dense partial sums moves.
backward  table-wise : the kernel that PRODUCES the gradient of the pooled rows (the dot interaction's
                       backward, b200_interaction_backward_scatter) stores every table-wise feature's rows
                       straight into the owning rank's [B_global, D_owned] buffer: the reverse all-to-all
                       happens inside the producer, and the fused optimizer kernel then reads local HBM.
                       A model whose gradient arrives some other way takes the pull path below.
          row-wise   : b200_tbe_backward_peer loads the gradient row of sample b from rank
          (and pull)   (b // B_local)'s gradient buffer while it streams the sorted ids, applies 1/W and
                       the fused optimizer (the all-gather, restricted to the rows this rank owns: with one
                       id per bag 1/W of the all-gather's traffic).

Three stream-ordered cross-rank barriers per step bracket the peer accesses; no pooled
[B_global, D_local] tensor, no NCCL pass over the embeddings and no split/cat copy exist.
Enabled for CUDA process groups with world_size > 1 (B200_PEER_DIST=0 disables it), for unweighted and weighted
collections and for 16-bit tables with stochastic rounding; gradient clipping, variable batch sizes and sequence
embeddings take the NCCL path in comm_ops.py / dist_data.py, which stays as the portable implementation and is what
the gloo tests run.
"""
from __future__ import annotations

import os
import warnings
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from b200_sparse import ops as _b200
from torchrec.distributed import comm_ops


def peer_dist_wanted(pg: Optional[dist.ProcessGroup], device: Optional[torch.device]) -> bool:
    if os.environ.get("B200_PEER_DIST", "1") == "0":
        return False
    if pg is None or device is None or device.type != "cuda":
        return False
    return dist.get_world_size(pg) > 1 and dist.get_backend(pg) == "nccl"


class PeerExchange:
    """Symmetric buffers of one sharded module for a given local batch size."""

    def __init__(self, pg: dist.ProcessGroup, device: torch.device, B_local: int, total_D: int, rw_dim: int,
                 feat_cols: Optional[List[Tuple[int, int]]] = None,
                 tw_map: Optional[List[Optional[Tuple[int, int]]]] = None, tw_pitch: int = 0,
                 rw_features: int = 0) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        self.W, self.rank = dist.get_world_size(pg), dist.get_rank(pg)
        self.B_local, self.total_D, self.rw_dim = B_local, total_D, rw_dim
        gname = pg.group_name
        try:
            symm_mem.enable_symm_mem_for_group(gname)
        except Exception:  # newer torch: rendezvous does it
            pass

        def make(shape, dtype=torch.float32) -> Tuple[torch.Tensor, object, torch.Tensor]:
            t = symm_mem.empty(shape, dtype=dtype, device=device)
            h = symm_mem.rendezvous(t, gname)
            ptrs = [int(p) for p in h.buffer_ptrs]
            return t, h, ptrs

        self.out, self._h_out, out_ptrs = make((B_local, total_D))
        self.grad, self._h_grad, grad_ptrs = make((B_local, total_D))
        # Gradient push (table-wise features): gin[r * B_local + b, local col] on the OWNER holds the
        # gradient of rank r's sample b.  feat_base / feat_pitch tell the producer where every feature's
        # rows go: table-wise ones into the owner's gin, the rest into my own `grad` (pulled by the owners).
        self.gin = self.gin_base = None
        self.tw_pitch = tw_pitch
        feat_base = feat_pitch = None
        self.pushed_steps = 0
        same_dim = feat_cols is not None and len({d for _c, d in feat_cols}) == 1
        if tw_map is not None and tw_pitch > 0 and same_dim and os.environ.get("B200_PEER_PUSH", "1") != "0":
            self.gin, self._h_gin, gin_ptrs = make((self.W * B_local, tw_pitch))
            self.gin_base = torch.tensor([gin_ptrs[self.rank]], dtype=torch.int64, device=device)
            fb, fp = [], []
            for (col, _dim), tw in zip(feat_cols, tw_map):
                if tw is None:
                    fb.append(grad_ptrs[self.rank] + 4 * col)
                    fp.append(total_D)
                else:
                    owner, local_col = tw
                    fb.append(gin_ptrs[owner] + 4 * (self.rank * B_local * tw_pitch + local_col))
                    fp.append(tw_pitch)
            feat_base = torch.tensor(fb, dtype=torch.int64, device=device)
            feat_pitch = torch.tensor(fp, dtype=torch.int64, device=device)
        # the dot-interaction backward writes its grad_sparse here (and, feature by feature, to the owners)
        self.sink = _b200.register_grad_sink(self.grad, feat_base, feat_pitch, self.out)
        self.out_bases = torch.tensor(out_ptrs, dtype=torch.int64, device=device)
        self.grad_bases = torch.tensor(grad_ptrs, dtype=torch.int64, device=device)
        self.stage = None
        self.stage_bases = None
        if rw_dim > 0:
            self.stage, self._h_stage, st_ptrs = make((self.W, B_local, rw_dim))
            # kept all-zero between steps: peers store only their non-empty partial bags and
            # b200_peer_stage_reduce re-zeroes what it consumed (the first forward's barrier A orders this)
            self.stage.zero_()
            slot = self.rank * B_local * rw_dim * 4  # my slot in every destination's staging buffer
            self.stage_bases = torch.tensor([p + slot for p in st_ptrs], dtype=torch.int64, device=device)
        # presence flags [B_local, F_rw, W]: peers set flags[b][f][their rank] next to a stored partial row and
        # the reduction reads only flagged slots (B200_PEER_FLAGS=0: the dense reduce over all W slots)
        self.flags = self.flag_bases = None
        self.rw_features = rw_features
        if rw_dim > 0 and rw_features > 0 and self.W <= 32 and os.environ.get("B200_PEER_FLAGS", "1") != "0":
            self.flags, self._h_flags, fl_ptrs = make((B_local, rw_features, self.W), torch.uint8)
            self.flags.zero_()
            self.flag_bases = torch.tensor(fl_ptrs, dtype=torch.int64, device=device)

        # True after a backward pass: its barrier C already proves that every rank is done with the
        # buffers of that step, so the next forward may skip barrier A (see _PeerLookup.forward)
        self.quiesced = False
        # Aliasing contract: `out` is ONE persistent buffer that the lookup kernels of every rank rewrite
        # through raw pointers (no autograd version bump).  A training forward therefore hands out a view of
        # it that stays valid until that step's backward has run; `outstanding` marks the window, and a
        # second forward inside it is refused instead of silently overwriting saved activations.
        # Forwards without autograd (evaluation) return a private copy.
        self.outstanding = False
        # B200_PEER_BARRIER=b200: our own barrier kernel (b200_peer_barrier: one 32-thread launch through the C
        # ABI with a device-resident epoch, usable inside a replayed CUDA graph).  Measured at N = 2 it is SLOWER
        # in the step than the symmetric-memory handle's barrier (5.89 vs 5.48 ms/step: its system-scope fences
        # wait for the remote stores to drain inside the barrier kernel), so the handle's barrier is the default.
        self.pad_bases = self.epoch_ctr = None
        if os.environ.get("B200_PEER_BARRIER", "torch") == "b200":
            self.pad, self._h_pad, pad_ptrs = make((max(self.W, 32),), torch.int32)
            self.pad.zero_()
            self.epoch_ctr = torch.zeros(1, dtype=torch.int32, device=device)
            torch.cuda.current_stream(device).synchronize()
            self._h_pad.barrier()  # every rank's pad is zeroed before anyone signals into it
            self.pad_bases = torch.tensor(pad_ptrs, dtype=torch.int64, device=device)

    def __del__(self) -> None:
        try:
            _b200.unregister_grad_sink(self.sink)
        except Exception:
            pass

    def barrier(self) -> None:
        if self.pad_bases is not None:
            _b200.peer_barrier(self.pad_bases, self.rank, self.W, self.epoch_ctr)
        else:
            self._h_out.barrier()


class _Part:
    """One table-batched lookup module with its column maps."""

    __slots__ = ("tbe", "row_wise", "dst_col", "src_col", "n_features", "gin_col", "flag_col")

    def __init__(self, tbe, row_wise: bool, dst_col: torch.Tensor, src_col: torch.Tensor, n_features: int,
                 gin_col: Optional[torch.Tensor] = None) -> None:
        self.tbe, self.row_wise, self.dst_col, self.src_col, self.n_features = tbe, row_wise, dst_col, src_col, n_features
        self.gin_col = gin_col  # table-wise: this module's feature columns inside my gin buffer
        self.flag_col: Optional[torch.Tensor] = None  # row-wise: f * W + my rank for each of its features


class PeerLookupDist:
    """compute_and_output_dist of a ShardedEmbeddingBagCollection through peer memory."""

    def __init__(self, sebc, pg: dist.ProcessGroup, device: torch.device) -> None:
        from torchrec.distributed.sharding.rw_sharding import BaseRwEmbeddingSharding
        from torchrec.distributed.sharding.tw_sharding import BaseTwEmbeddingSharding
        self._pg, self._device = pg, device
        self._W, self._rank = dist.get_world_size(pg), dist.get_rank(pg)
        self._exchanges: Dict[int, PeerExchange] = {}
        self._parts: List[List[_Part]] = []  # per sharding, per module
        # Output columns follow the MODEL's feature order (the unsharded EmbeddingBagCollection's
        # KeyedTensor), not the per-sharding-type grouping of the NCCL path: the kernels place every
        # feature where the consumer wants it, so no reorder copy follows the exchange.
        uses: Dict[str, int] = {}
        for cfg in sebc._embedding_bag_configs:
            for f in cfg.feature_names:
                uses[f] = uses.get(f, 0) + 1
        self.embedding_names: List[str] = []
        self.embedding_dims: List[int] = []
        col_of: Dict[str, int] = {}
        col = 0
        for cfg in sebc._embedding_bag_configs:
            for f in cfg.feature_names:
                name = f + "@" + cfg.name if uses[f] > 1 else f
                col_of[name] = col
                self.embedding_names.append(name)
                self.embedding_dims.append(cfg.embedding_dim)
                col += cfg.embedding_dim
        self._total_D = col
        self._tw_of: Dict[str, Tuple[int, int]] = {}  # embedding name -> (owner rank, column in its gin)
        self._tw_pitch = 0
        self._rw_dim = 0
        self._weighted = False  # the collection's features carry per-id weights (id_score_list features)
        self._rw_col4: Optional[torch.Tensor] = None
        self._rw_feat_col4: Optional[torch.Tensor] = None
        self._rw_features = 0

        def cols(names: List[str]) -> torch.Tensor:
            return torch.tensor([col_of[n] for n in names], dtype=torch.int32, device=device)

        for sharding, lookup in zip(sebc._sharding_type_to_sharding.values(), sebc._lookups):
            # a collection is either unweighted (id_list features) or weighted (id_score_list features)
            if len(lookup._emb_modules) and len(lookup._score_emb_modules):
                raise NotImplementedError("mixed weighted / unweighted lookups take the NCCL path")
            self._weighted = bool(len(lookup._score_emb_modules)) or getattr(self, "_weighted", False)
            mods = list(lookup._score_emb_modules) if len(lookup._score_emb_modules) else list(lookup._emb_modules)
            if any(m.emb_module.optimizer_args.gradient_clipping for m in mods):
                raise NotImplementedError("gradient clipping takes the NCCL path")
            parts: List[_Part] = []
            if isinstance(sharding, BaseTwEmbeddingSharding):
                # every rank's table-wise features, in that rank's module order: the layout of its gin
                tw_dims = []
                per_rank = (sharding._score_grouped_embedding_configs_per_rank if self._weighted
                            else sharding._grouped_embedding_configs_per_rank)
                for r, cfgs in enumerate(per_rank):
                    local = 0
                    for cfg in cfgs:
                        for n, d in zip(cfg.embedding_names(), cfg.embedding_dims()):
                            self._tw_of[n] = (r, local)
                            local += d
                    tw_dims.append(local)
                self._tw_pitch = max(tw_dims) if tw_dims else 0
                base = 0
                for m in mods:
                    c = cols(m.config.embedding_names())
                    gin_col = (m.emb_module.D_offsets[:-1] + base).to(torch.int32)
                    parts.append(_Part(m.emb_module, False, c, c, m.config.num_features(), gin_col))
                    base += m.emb_module.total_D
            elif isinstance(sharding, BaseRwEmbeddingSharding):
                if self._rw_dim:
                    raise NotImplementedError("one row-wise sharding per module")
                local0 = 0
                chunk_cols: List[int] = []
                feat_col4: List[int] = [0]
                # the feature numbering of the presence flags must be the same on every rank: the row-wise
                # features of ALL ranks' modules are the same list (every rank holds a row block of each table)
                for m in mods:
                    stage_cols = (m.emb_module.D_offsets[:-1] + local0).to(torch.int32)
                    names = m.config.embedding_names()
                    part = _Part(m.emb_module, True, stage_cols, cols(names), m.config.num_features())
                    f0 = len(feat_col4) - 1
                    part.flag_col = torch.tensor([(f0 + k) * self._W + self._rank for k in range(len(names))],
                                                 dtype=torch.int32, device=device)
                    parts.append(part)
                    for n, d in zip(names, m.config.embedding_dims()):
                        chunk_cols.extend(col_of[n] + 4 * k for k in range(d // 4))
                        feat_col4.append(feat_col4[-1] + d // 4)
                    local0 += m.emb_module.total_D
                self._rw_dim = local0
                self._rw_col4 = torch.tensor(chunk_cols, dtype=torch.int32, device=device)
                self._rw_feat_col4 = torch.tensor(feat_col4, dtype=torch.int32, device=device)
                self._rw_features = len(feat_col4) - 1
            else:
                raise NotImplementedError(type(sharding).__name__)
            self._parts.append(parts)
        if any(d % 4 for d in self.embedding_dims):
            raise NotImplementedError("embedding dims must be multiples of 4")
        # every lookup module's placeholder parameter enters the autograd graph, as in the per-module
        # path: DDP registers them and waits for each one before it reduces a bucket
        self._placeholders = [p.tbe.placeholder_autograd_tensor for parts in self._parts for p in parts]
        if not self._placeholders:  # a rank without tables still joins every barrier, forward and backward
            self._placeholders = [torch.zeros(0, device=device, requires_grad=True)]

    def exchange(self, B_local: int) -> PeerExchange:
        px = self._exchanges.get(B_local)
        if px is None:
            feat_cols = []
            c0 = 0
            for d in self.embedding_dims:
                feat_cols.append((c0, d))
                c0 += d
            tw_map = [self._tw_of.get(n) for n in self.embedding_names]
            px = PeerExchange(self._pg, self._device, B_local, self._total_D, self._rw_dim, feat_cols, tw_map,
                              self._tw_pitch, self._rw_features)
            self._exchanges[B_local] = px
        return px

    def __call__(self, dist_input, B_local: int) -> torch.Tensor:
        """dist_input: SparseFeaturesList (one per sharding) after the KJT all-to-all; B_local: this
        rank's batch size (every rank must use the same one: no variable batch)."""
        work = []  # (part, indices, offsets, per-sample weights)
        for parts, sf in zip(self._parts, dist_input):
            kjt = sf.id_score_list_features if self._weighted else sf.id_list_features
            if not parts:
                continue
            pieces = kjt.split([p.n_features for p in parts]) if len(parts) > 1 else [kjt]
            for p, piece in zip(parts, pieces):
                assert piece.stride() == B_local * self._W, "variable batch size per rank is out of scope"
                idx, off = p.tbe.prepare_indices(piece.values(), piece.offsets())
                psw = piece.weights().float().contiguous() if self._weighted else None
                work.append((p, idx, off, psw))
        px = self.exchange(B_local)
        return _PeerLookup.apply(self, px, work, *self._placeholders)


class _PeerLookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, owner: PeerLookupDist, px: PeerExchange, work, *placeholders):
        from fbgemm_gpu import _tbe_autograd as ta
        need_grad = any(ctx.needs_input_grad[3:])
        ctx.n_placeholders = len(placeholders)
        if px.outstanding and os.environ.get("B200_PEER_ALLOW_REENTRY", "0") != "1":
            raise RuntimeError(
                "peer-memory embedding lookup re-entered before the backward of its previous training forward: "
                "the pooled output lives in one persistent symmetric buffer that this call would overwrite "
                "(see PeerExchange.outstanding). Run backward first, call the module under torch.no_grad() for "
                "evaluation, or set B200_PEER_DIST=0 to use the NCCL path, which returns fresh tensors.")
        main = torch.cuda.current_stream(px.out.device)
        sorts = []
        for p, idx, off, psw in work:
            m = p.tbe
            T = m.D_offsets.numel() - 1
            B = (off.numel() - 1) // T
            if need_grad and idx.numel() > 0:
                side = ta._side_stream(idx.device)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    s = torch.ops.fbgemm.b200_tbe_sort_indices(idx, off, m.hash_size_cumsum,
                                                               m.total_hash_size_bits, B, psw is not None,
                                                               getattr(m, "max_table_bits", 0))
                    ev = torch.cuda.Event()
                    ev.record(side)
                for t in (idx, off):
                    t.record_stream(side)
                for t in s:
                    t.record_stream(main)
                sorts.append((s, ev, B))
            else:
                sorts.append((None, None, B))
        # A: every rank must be done with the buffers of the previous use.  After a training step that is
        # already known: barrier C of its backward comes after each rank's last read of `out` (the
        # interaction backward) and of the staging buffer, and peers only overwrite `grad` after barrier
        # B below.  Only forward-after-forward (evaluation) needs the explicit barrier.  Every rank runs
        # the same sequence, so the decision is identical everywhere.
        if not px.quiesced:
            px.barrier()
        px.quiesced = False
        for (p, idx, off, psw) in work:
            m = p.tbe
            bases = px.stage_bases if p.row_wise else px.out_bases
            pitch = px.rw_dim if p.row_wise else px.total_D
            _b200.tbe_forward_peer(m.weights_dev, m.weights_offsets, m.D_offsets, m.total_D, m.max_D,
                                   m._all_D_mult8, idx, off, psw, int(m.pooling_mode), bases, px.B_local, pitch,
                                   p.dst_col, skip_empty=p.row_wise,
                                   flag_bases=px.flag_bases if p.row_wise else None,
                                   flag_pitch=px.rw_features * px.W, flag_col=p.flag_col if p.row_wise else None)
        px.barrier()  # B: all peer stores have landed
        if px.stage is not None:
            if px.flags is not None:
                _b200.peer_stage_reduce_flagged(px.stage, px.flags, px.W, px.B_local, px.rw_dim,
                                                owner._rw_feat_col4, px.out, owner._rw_col4)
            else:
                _b200.peer_stage_reduce(px.stage, px.W, px.B_local, px.rw_dim, px.out, owner._rw_col4,
                                        clear_stage=True)
        ctx.px, ctx.work, ctx.sorts = px, work, sorts
        if not need_grad:
            return px.out.clone()  # evaluation: the caller may keep it across further forwards
        px.outstanding = True
        return px.out

    @staticmethod
    def backward(ctx, grad_out):
        from fbgemm_gpu import _tbe_autograd as ta
        px = ctx.px
        px.outstanding = False
        pushed = px.sink.scattered
        px.sink.scattered = False
        if grad_out.data_ptr() != px.grad.data_ptr() or not grad_out.is_contiguous():
            if pushed:
                raise RuntimeError(
                    "the pooled embeddings' gradient was scattered to the owning ranks by its producer but autograd "
                    "delivered a different tensor (the embeddings feed more than one consumer); set "
                    "B200_PEER_PUSH=0 to use the pull exchange")
            px.grad.copy_(grad_out)  # producers that know the sink (ops.register_grad_sink) skip this pass
        px.pushed_steps += int(pushed)
        px.barrier()  # C: every rank's gradient buffer is complete / every pushed row has landed
        px.quiesced = True
        scale = 1.0 / px.W if comm_ops.GRADIENT_DIVISION else 1.0
        for (p, idx, off, psw), (s, ev, B) in zip(ctx.work, ctx.sorts):
            if s is None:
                continue
            m = p.tbe
            a = m.optimizer_args
            assert not a.gradient_clipping  # refused at build time (PeerLookupDist.__init__)
            if ev is not None:
                torch.cuda.current_stream(idx.device).wait_event(ev)
            keys, payload, bag = s
            srt = _b200.SortedIndices(keys, payload, bag if psw is not None else None, keys.numel(),
                                      max(1, m.total_hash_size_bits), None)
            if pushed and not p.row_wise:  # my features' gradients of all B_global samples are in my gin
                bases, B_src, pitch, src_col = px.gin_base, px.W * px.B_local, px.tw_pitch, p.gin_col
            else:
                bases, B_src, pitch, src_col = px.grad_bases, px.B_local, px.total_D, p.src_col
            _b200.tbe_backward_peer(bases, B_src, pitch, src_col, scale, srt, m.weights_dev,
                                    m.weights_offsets, m.D_offsets, m.hash_size_cumsum, B, m.max_D, m._all_D_mult8,
                                    off, psw, int(m.pooling_mode), ta._OPT_CODE[m.optimizer],
                                    m.momentum1_dev if m.momentum1_dev.numel() else None, m.momentum1_offsets,
                                    a.learning_rate, a.eps, *m._next_sr_state())
        ctx.work = ctx.sorts = None
        return (None, None, None) + (None,) * ctx.n_placeholders


def try_build(sebc, pg, device) -> Optional[PeerLookupDist]:
    if not peer_dist_wanted(pg, device):
        return None
    try:
        return PeerLookupDist(sebc, pg, device)
    except NotImplementedError as e:
        warnings.warn(f"peer-memory embedding exchange not used ({e}); falling back to NCCL collectives")
        return None
