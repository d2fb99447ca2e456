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
from typing import Callable, Deque, Generic, Iterator, List, Optional, Tuple, TypeVar

import torch
from torch import nn

from torchrec.distributed.types import ShardedModule
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

In = TypeVar("In")
Out = TypeVar("Out")


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
                 loss_log_path: Optional[str] = None) -> None:
        super().__init__(model, optimizer, device)
        self._loss_log_path = loss_log_path
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
        losses, output = self._model(self._batch_i)
        if self._model.training:
            if self._loss_log_path is not None and _rank() == 0:
                with open(self._loss_log_path, "a") as f:  # same line format as the fork's log
                    f.write(str(losses.detach().cpu().numpy().tolist()) + "\n")
            torch.sum(losses, dim=0).backward()
            self._optimizer.step()
        # ... phase 2: the host read of the receive sizes happens once the whole step of batch i is
        # queued, when that exchange has long finished, so neither the host nor the GPU waits.  (The
        # reference does both phases between forward and backward, :529: the GPU idles during the sync.)
        if self._batch_ip1 is not None:
            self._complete_sparse_data_dist()
        self._batch_i, self._batch_ip1 = self._batch_ip1, self._batch_ip2
        return output
