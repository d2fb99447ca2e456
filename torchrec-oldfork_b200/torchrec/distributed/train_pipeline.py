"""TrainPipelineSparseDist: the caller of the sparse hot path (reference
torchrec/distributed/train_pipeline.py:422-557).

Three batches are in flight: batch i+2 is copied host -> device on a memcpy stream, batch i+1 goes
through the sparse input distribution (bucketize / permute kernels, the KJT all-to-alls and the ONE host
read of the length table) on a data-dist stream, and batch i runs forward / backward / optimizer on the
current stream.  The host therefore blocks on the lengths of batch i+1 while the GPU is busy with
batch i, instead of draining the queue in the middle of every step.

The reference finds the sharded modules' inputs by FX-tracing the model; here every ShardedModule takes
`sparse_features_fn(batch)` (default: `batch.sparse_features`), which covers DLRM-style models whose
sharded collections all consume the batch's KeyedJaggedTensor.  The reference fork appends every step's
losses to a hard-coded file on rank 0 (:538-543); here that is the optional `loss_log_path` argument
(off by default: it costs a device -> host read per step).
"""
from __future__ import annotations

from collections import deque
from typing import Callable, Deque, Generic, Iterator, List, Optional, TypeVar

import torch
from torch import nn

from torchrec.distributed.types import ShardedModule
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

In = TypeVar("In")
Out = TypeVar("Out")


# ---------------------------------------------------------------------------------------------
# CUDA-graph replay of the steady-state step (SURVEY §8 f1 "then CUDA-graph the steady-state step").
#
# What is captured: forward and backward (with the fused table update and the peer-memory exchange) of ONE
# batch whose sparse input distribution has already run, ending in one flat buffer of dense gradients.  What
# stays eager: the all-reduce of that buffer and the dense optimizer step (a handful of launches), and, on
# their own streams, the host->device copy of batch i+2 and the input distribution of batch i+1 (its
# all-to-alls need the host to read the receive sizes).  Every step the distributed ids of batch i are
# copied into the graph's fixed input buffers and the graph is replayed: ~300 launches, the cross-rank
# barriers and the autograd bookkeeping of a step cost the host one cudaGraphLaunch.
#
# A graph is only valid for the exact shapes it was captured with - including the per-key id counts, which
# become pointer offsets inside it - and for the learning rates, which are kernel arguments.  The signature
# below covers all of them; a step whose signature has no graph runs eagerly (and is captured if the
# signature repeats), so ragged inputs keep working.
def _sig(obj):
    from torchrec.distributed.embedding_types import SparseFeatures, SparseFeaturesList
    if obj is None:
        return None
    if isinstance(obj, torch.Tensor):
        return (tuple(obj.shape), obj.dtype)
    if isinstance(obj, KeyedJaggedTensor):
        return ("kjt", tuple(obj.keys()), obj.stride(), tuple(obj.length_per_key()),
                tuple(_sig(t) for t in (obj._values, obj._weights, obj._lengths, obj._offsets)))
    if isinstance(obj, SparseFeatures):
        return ("sf", _sig(obj.id_list_features), _sig(obj.id_score_list_features))
    if isinstance(obj, SparseFeaturesList):
        return ("sfl",) + tuple(_sig(f) for f in obj.features)
    if isinstance(obj, (list, tuple)):
        return tuple(_sig(o) for o in obj)
    if hasattr(obj, "dense_features") and hasattr(obj, "sparse_features") and hasattr(obj, "labels"):
        return ("batch", _sig(obj.dense_features), _sig(obj.sparse_features), _sig(obj.labels))
    raise TypeError(f"CUDA-graph mode does not know how to hold a {type(obj).__name__} at a fixed address")


def _static_copy(obj):
    """Same structure, every tensor cloned (these clones are the graph's input buffers)."""
    from torchrec.distributed.embedding_types import SparseFeatures, SparseFeaturesList
    if obj is None:
        return None
    if isinstance(obj, torch.Tensor):
        return obj.clone()
    if isinstance(obj, KeyedJaggedTensor):
        c = lambda t: None if t is None else t.clone()  # noqa: E731
        return KeyedJaggedTensor(keys=list(obj.keys()), values=c(obj._values), weights=c(obj._weights),
                                 lengths=c(obj._lengths), offsets=c(obj._offsets), stride=obj.stride(),
                                 length_per_key=list(obj.length_per_key()))
    if isinstance(obj, SparseFeatures):
        return SparseFeatures(id_list_features=_static_copy(obj.id_list_features),
                              id_score_list_features=_static_copy(obj.id_score_list_features))
    if isinstance(obj, SparseFeaturesList):
        return SparseFeaturesList([_static_copy(f) for f in obj.features])
    if isinstance(obj, (list, tuple)):
        return type(obj)(_static_copy(o) for o in obj)
    return type(obj)(dense_features=_static_copy(obj.dense_features),
                     sparse_features=_static_copy(obj.sparse_features), labels=_static_copy(obj.labels))


def _tensors(obj, out: list) -> list:
    from torchrec.distributed.embedding_types import SparseFeatures, SparseFeaturesList
    if obj is None:
        return out
    if isinstance(obj, torch.Tensor):
        out.append(obj)
    elif isinstance(obj, KeyedJaggedTensor):
        out.extend(t for t in (obj._values, obj._weights, obj._lengths, obj._offsets) if t is not None)
    elif isinstance(obj, SparseFeatures):
        _tensors(obj.id_list_features, out)
        _tensors(obj.id_score_list_features, out)
    elif isinstance(obj, SparseFeaturesList):
        for f in obj.features:
            _tensors(f, out)
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            _tensors(o, out)
    else:
        _tensors(obj.dense_features, out)
        _tensors(obj.sparse_features, out)
        _tensors(obj.labels, out)
    return out


class _StepGraph:
    __slots__ = ("graph", "static", "slots", "output", "launches", "flat", "group", "grads")


def _rank() -> int:
    import torch.distributed as dist
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def _to_device(batch, device: torch.device, non_blocking: bool):
    return batch.to(device=device, non_blocking=non_blocking)


class TrainPipelineBase(Generic[In, Out]):
    """Two-stage pipeline (reference :75-160): H2D copy of batch i+1 overlaps the step of batch i."""

    def __init__(self, model: nn.Module, optimizer: torch.optim.Optimizer, device: torch.device) -> None:
        self._model, self._optimizer, self._device = model, optimizer, device
        self._memcpy_stream = torch.cuda.Stream(device=device, priority=-1) if device.type == "cuda" else None
        self._cur_batch = None
        self._connected = False

    def _connect(self, it: Iterator[In]) -> None:
        with torch.cuda.stream(self._memcpy_stream):
            self._cur_batch = _to_device(next(it), self._device, True)
        self._connected = True

    def progress(self, dataloader_iter: Iterator[In]) -> Out:
        if not self._connected:
            self._connect(dataloader_iter)
        if self._model.training:
            self._optimizer.zero_grad()
        cur = self._cur_batch
        torch.cuda.current_stream(self._device).wait_stream(self._memcpy_stream)
        cur.record_stream(torch.cuda.current_stream(self._device))
        with torch.cuda.stream(self._memcpy_stream):
            try:
                self._cur_batch = _to_device(next(dataloader_iter), self._device, True)
            except StopIteration:
                self._cur_batch = None
                self._connected = False
        losses, output = self._model(cur)
        if self._model.training:
            torch.sum(losses, dim=0).backward()
            self._optimizer.step()
        return output


class TrainPipelineSparseDist(TrainPipelineBase[In, Out]):
    def __init__(self, model: nn.Module, optimizer: torch.optim.Optimizer, device: torch.device,
                 sparse_features_fn: Optional[Callable[[In], KeyedJaggedTensor]] = None,
                 loss_log_path: Optional[str] = None, cuda_graph: bool = False, graph_warmup: int = 11) -> None:
        super().__init__(model, optimizer, device)
        self._loss_log_path = loss_log_path
        # cuda_graph: replay the step as one CUDA graph once `graph_warmup` eager steps have run (DDP needs
        # 11 eager iterations before its backward may be captured) and the step's signature repeats
        # B200_PIPE_COMPLETE=early: read the receive sizes and start the values all-to-alls of batch i+1 between
        # forward and backward of batch i (as the reference, :529), which puts them AHEAD of this step's gradient
        # all-reduce in the NCCL queue (default; measured 5.27 vs 5.33 ms/step at N = 2); "late": after the whole step
        # has been queued
        import os
        self._complete_early = os.environ.get("B200_PIPE_COMPLETE", "early") == "early"
        self._cuda_graph = bool(cuda_graph) and loss_log_path is None
        self._graph_warmup = graph_warmup
        self._eager_steps = 0
        self._graphs: dict = {}
        self._last_sig = None
        self.graph_replays = 0
        if device.type != "cuda":
            raise RuntimeError("TrainPipelineSparseDist overlaps CUDA streams; use TrainPipelineBase on CPU")
        # high priority: the small bucketize / permute / reduction kernels whose results the host waits for
        # must not queue behind the thousands of CTAs of the previous batch's backward
        self._data_dist_stream = torch.cuda.Stream(device=device, priority=-1)
        self._features_of = sparse_features_fn or (lambda b: b.sparse_features)
        self._sharded: List[ShardedModule] = [m for m in model.modules() if isinstance(m, ShardedModule)]
        for m in self._sharded:
            m._pipelined_inputs = deque()
        self._batch_i = self._batch_ip1 = self._batch_ip2 = None
        self._dist_done: Deque[torch.cuda.Event] = deque()
        self._issued: list = []

    # reference _start_data_dist (:254-300) + _wait_sparse_data_dist, split in two so that the host sync
    # of phase 1 is taken only after the whole step of the previous batch has been queued
    def _issue_sparse_data_dist(self, batch) -> None:
        """Enqueues bucketize / permute and the lengths all-to-alls of `batch` on the data-dist stream."""
        with torch.cuda.stream(self._data_dist_stream):
            self._data_dist_stream.wait_stream(self._memcpy_stream)
            batch.record_stream(self._data_dist_stream)
            kjt = self._features_of(batch)
            self._issued = []
            for m in self._sharded:
                ctx = m.create_context()
                self._issued.append((m, ctx, m.input_dist(ctx, kjt)))

    def _complete_sparse_data_dist(self) -> None:
        """Reads the receive sizes, runs the values all-to-alls and the recat on the data-dist stream and
        leaves the distributed features in each module's pipelined queue."""
        main = torch.cuda.current_stream(self._device)
        with torch.cuda.stream(self._data_dist_stream):
            done = list(self._issued)
            for m, ctx, awaitable in self._issued:
                dist_input = awaitable.wait()
                record = getattr(dist_input, "record_stream", None)
                if record is not None:
                    record(main)
                m._pipelined_inputs.append((ctx, dist_input))
            self._issued = []
            ev = torch.cuda.Event()
            ev.record(self._data_dist_stream)
            self._dist_done.append(ev)
            for _m, ctx, _a in done:
                ctx.inputs_ready_event = ev  # lookups on their own stream wait for this, not for `main`

    def _start_sparse_data_dist(self, batch) -> None:
        self._issue_sparse_data_dist(batch)
        self._complete_sparse_data_dist()

    def reset(self) -> None:
        """Drops the prefetched batches and their queued sparse distributions."""
        torch.cuda.current_stream(self._device).wait_stream(self._data_dist_stream)
        for m in self._sharded:
            m._pipelined_inputs.clear()
        self._dist_done.clear()
        self._issued = []
        self._batch_i = self._batch_ip1 = self._batch_ip2 = None
        self._connected = False

    # ---- CUDA-graph replay
    def _lr_signature(self):
        lrs = [g.get("lr") for g in self._optimizer.param_groups]
        for m in self._model.modules():
            args = getattr(m, "optimizer_args", None)
            if args is not None and hasattr(args, "learning_rate"):
                lrs.append(float(args.learning_rate))
        return tuple(lrs)

    def _graphed_step(self):
        """Replays (capturing first if needed) the step of batch i; None = run this step eagerly."""
        if self._eager_steps < self._graph_warmup:
            return None
        if any(len(m._pipelined_inputs) == 0 for m in self._sharded):
            return None
        step_in = (self._batch_i, [m._pipelined_inputs[0][1] for m in self._sharded])
        try:
            sig = (_sig(step_in), self._lr_signature())
        except TypeError as e:
            import warnings
            warnings.warn(f"TrainPipelineSparseDist: CUDA-graph mode switched off ({e})")
            self._cuda_graph = False
            return None
        sg = self._graphs.get(sig)
        if sg is None:
            if sig != self._last_sig:  # capture only shapes that repeat
                self._last_sig = sig
                return None
            sg = self._capture(step_in)
            self._graphs[sig] = sg
        for dst, src in zip(sg.slots, _tensors(step_in, [])):
            dst.copy_(src, non_blocking=True)
        sg.graph.replay()
        for m in self._sharded:
            m._pipelined_inputs.popleft()
        # Eager tail: the data-parallel gradient reduction and the dense optimizer.  The all-reduce stays
        # OUT of the graph on purpose: a captured NCCL kernel replayed from the graph's own stream would run
        # unordered against the eager all-to-alls of the next batch's input distribution on the same
        # communicator (and a second communicator deadlocked against the peer barriers when tried).
        if sg.flat is not None:
            import torch.distributed as dist
            dist.all_reduce(sg.flat, group=sg.group)
            sg.flat.mul_(1.0 / dist.get_world_size(sg.group))
        for p, g in sg.grads:
            p.grad = g
        self._optimizer.step()
        self.graph_replays += 1
        from b200_sparse.ops import KernelProfile
        KernelProfile.launches += sg.launches
        return sg.output

    def _capture(self, step_in) -> "_StepGraph":
        """Captures forward + backward (fused table update included) of one step on buffers of its own; capture
        records without running, so the caller then copies the real inputs in and replays.  The dense
        gradients are packed into one flat buffer at the end of the graph; under DistributedDataParallel the
        capture runs in no_sync() mode (its reducer hooks are no-ops) and the caller reduces the flat buffer."""
        import contextlib
        import torch.distributed as dist
        from torch.nn.parallel import DistributedDataParallel
        from b200_sparse.ops import KernelProfile
        sg = _StepGraph()
        sg.static = _static_copy(step_in)
        sg.slots = _tensors(sg.static, [])  # before the capture fills the KJTs' lazy caches
        batch, dist_inputs = sg.static
        ctxs = [m._pipelined_inputs[0][0] for m in self._sharded]
        ddp = next((m for m in self._model.modules() if isinstance(m, DistributedDataParallel)), None)
        params = [p for g in self._optimizer.param_groups for p in g["params"]]
        was_profiling, KernelProfile.enabled = KernelProfile.enabled, False  # timing events cannot be captured
        n0 = KernelProfile.launches
        self._optimizer.zero_grad(set_to_none=True)
        torch.cuda.synchronize(self._device)
        sg.graph = torch.cuda.CUDAGraph()
        sg.flat, sg.group = None, None
        # thread_local: other threads of the process (NCCL watchdog, data loading) keep making CUDA calls that a
        # "global" capture would treat as errors; everything this step launches goes to capturing streams
        import os
        mode = os.environ.get("B200_GRAPH_CAPTURE_MODE", "thread_local")
        with torch.cuda.graph(sg.graph, capture_error_mode=mode), \
                (ddp.no_sync() if ddp is not None else contextlib.nullcontext()):
            start = torch.cuda.Event()
            start.record(torch.cuda.current_stream(self._device))
            for m, ctx, d in zip(self._sharded, ctxs, dist_inputs):
                ctx.inputs_ready_event = start  # an event of THIS capture: lookups fork from the graph's root
                m._pipelined_inputs.appendleft((ctx, d))
            losses, sg.output = self._model(batch)
            torch.sum(losses, dim=0).backward()
            with_grad = [p for p in params if p.grad is not None]
            if ddp is not None and with_grad:
                sg.flat = torch.cat([p.grad.reshape(-1) for p in with_grad])
        if sg.flat is not None:
            sg.group = ddp.process_group
            views, pos = [], 0
            for p in with_grad:
                views.append(sg.flat[pos: pos + p.numel()].view_as(p))
                pos += p.numel()
            sg.grads = list(zip(with_grad, views))
        else:
            sg.grads = [(p, p.grad) for p in with_grad]
        sg.launches = KernelProfile.launches - n0
        KernelProfile.launches = n0
        KernelProfile.enabled = was_profiling
        # the capture consumed the static entries it pushed; the real entries of batch i are still queued
        return sg

    def _copy(self, it: Iterator[In]):
        try:
            host = next(it)
        except StopIteration:
            return None
        kjt = self._features_of(host)
        if kjt is not None and not kjt.values().is_cuda:
            kjt.length_per_key()  # per-key totals from the HOST copy: the table-wise exchange then needs no
            #                       device read for its send sizes (they travel with the KJT through .to())
        with torch.cuda.stream(self._memcpy_stream):
            return _to_device(host, self._device, True)

    def _connect(self, it: Iterator[In]) -> None:
        self._batch_i = self._copy(it)
        if self._batch_i is None:
            raise StopIteration
        self._start_sparse_data_dist(self._batch_i)
        self._batch_ip1 = self._copy(it)
        self._connected = True

    def progress(self, dataloader_iter: Iterator[In]) -> Out:
        if not self._connected:
            self._connect(dataloader_iter)
        if self._batch_i is None:
            raise StopIteration
        if self._model.training:
            self._optimizer.zero_grad()
        main = torch.cuda.current_stream(self._device)
        # batch i: its sparse distribution was issued one step ago
        main.wait_event(self._dist_done.popleft())
        main.wait_stream(self._memcpy_stream)
        self._batch_i.record_stream(main)
        # batch i+2: host -> device
        self._batch_ip2 = self._copy(dataloader_iter)
        # batch i+1, phase 1: its lengths all-to-alls enter the NCCL queue ahead of this step's gradient
        # all-reduces ...
        if self._batch_ip1 is not None:
            self._issue_sparse_data_dist(self._batch_ip1)
        completed = False
        output = self._graphed_step() if self._cuda_graph and self._model.training else None
        if output is None:
            losses, output = self._model(self._batch_i)
            if self._complete_early and self._batch_ip1 is not None:
                self._complete_sparse_data_dist()
                completed = True
            if self._model.training:
                if self._loss_log_path is not None and _rank() == 0:
                    with open(self._loss_log_path, "a") as f:  # same line format as the fork's log
                        f.write(str(losses.detach().cpu().numpy().tolist()) + "\n")
                torch.sum(losses, dim=0).backward()
                self._optimizer.step()
            self._eager_steps += 1
        # ... phase 2: the host read of the receive sizes happens once the whole step of batch i is
        # queued, when that exchange has long finished, so neither the host nor the GPU waits.  (The
        # reference does both phases between forward and backward, :529: the GPU idles during the sync.)
        if self._batch_ip1 is not None and not completed:
            self._complete_sparse_data_dist()
        self._batch_i, self._batch_ip1 = self._batch_ip1, self._batch_ip2
        return output
