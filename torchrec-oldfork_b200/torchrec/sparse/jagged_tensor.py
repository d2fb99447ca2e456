"""JaggedTensor / KeyedJaggedTensor / KeyedTensor with the reference's API
(torchrec/sparse/jagged_tensor.py: JaggedTensor :153, KeyedJaggedTensor :614, KeyedTensor :1101).

Layout contract (unchanged): values are key-major then batch; lengths is [F * B] flattened
(key f, sample b at f * B + b); offsets = complete cumsum of lengths.  The integer work runs
on torch.ops.fbgemm.asynchronous_complete_cumsum / permute_2D_sparse_data (:36, :946), i.e.
on the sm_100a kernels.  `length_per_key` may be passed at construction (as the reference
allows, :659-695) so that split()/permute() never synchronise with the device.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch


def _offsets_of(lengths: torch.Tensor) -> torch.Tensor:
    return torch.ops.fbgemm.asynchronous_complete_cumsum(lengths)


def _lengths_of(offsets: torch.Tensor) -> torch.Tensor:
    return offsets[1:] - offsets[:-1]


def _prefix(xs: List[int]) -> List[int]:
    out, acc = [0], 0
    for x in xs:
        acc += x
        out.append(acc)
    return out


class _Jagged:
    """Shared values/weights/lengths/offsets storage with lazy lengths<->offsets."""

    def _init_jagged(self, values, weights, lengths, offsets) -> None:
        if lengths is None and offsets is None:
            raise AssertionError("Must provide lengths or offsets")
        self._values: torch.Tensor = values
        self._weights: Optional[torch.Tensor] = weights
        self._lengths: Optional[torch.Tensor] = lengths
        self._offsets: Optional[torch.Tensor] = offsets

    def lengths(self) -> torch.Tensor:
        if self._lengths is None:
            self._lengths = _lengths_of(self._offsets)
        return self._lengths

    def offsets(self) -> torch.Tensor:
        if self._offsets is None:
            self._offsets = _offsets_of(self._lengths)
        return self._offsets

    def values(self) -> torch.Tensor:
        return self._values

    def weights(self) -> torch.Tensor:
        if self._weights is None:
            raise AssertionError("This (Keyed)JaggedTensor doesn't have weights.")
        return self._weights

    def weights_or_none(self) -> Optional[torch.Tensor]:
        return self._weights

    def _tensors(self) -> List[torch.Tensor]:
        return [t for t in (self._values, self._weights, self._lengths, self._offsets) if t is not None]

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        for t in self._tensors():
            t.record_stream(stream)


class JaggedTensor(_Jagged):
    def __init__(self, values: torch.Tensor, weights: Optional[torch.Tensor] = None,
                 lengths: Optional[torch.Tensor] = None, offsets: Optional[torch.Tensor] = None) -> None:
        self._init_jagged(values, weights, lengths, offsets)

    @staticmethod
    def empty(is_weighted: bool = False) -> "JaggedTensor":
        return JaggedTensor(values=torch.tensor([]), weights=torch.tensor([]) if is_weighted else None,
                            lengths=torch.tensor([], dtype=torch.int32), offsets=torch.tensor([], dtype=torch.int32))

    @staticmethod
    def from_dense_lengths(values: torch.Tensor, lengths: torch.Tensor,
                           weights: Optional[torch.Tensor] = None) -> "JaggedTensor":
        mask = torch.arange(values.size(1), device=values.device).unsqueeze(0) < lengths.unsqueeze(1)
        return JaggedTensor(values=values[mask], weights=weights[mask] if weights is not None else None,
                            lengths=lengths)

    @staticmethod
    def from_dense(values: List[torch.Tensor], weights: Optional[List[torch.Tensor]] = None) -> "JaggedTensor":
        lengths = torch.tensor([v.numel() for v in values], dtype=torch.int32)
        return JaggedTensor(values=torch.cat(values), weights=torch.cat(weights) if weights is not None else None,
                            lengths=lengths, offsets=torch.tensor(_prefix(lengths.tolist()), dtype=torch.int32))

    def to_dense(self) -> List[torch.Tensor]:
        bounds = self.offsets().tolist()
        return [self._values[bounds[i]:bounds[i + 1]] for i in range(len(bounds) - 1)]

    def to_padded_dense(self, desired_length: Optional[int] = None, padding_value: float = 0.0) -> torch.Tensor:
        lengths = self.lengths()
        n = int(lengths.numel())
        width = int(lengths.max().item()) if desired_length is None else desired_length
        out = self._values.new_full((n, width), padding_value)
        cols = torch.arange(width, device=lengths.device).unsqueeze(0)
        mask = cols < lengths.unsqueeze(1).clamp(max=width)
        src_pos = (self.offsets()[:-1].unsqueeze(1) + cols)[mask]
        out[mask] = self._values[src_pos.long()]
        return out

    def to(self, device: torch.device, non_blocking: bool = False) -> "JaggedTensor":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)  # noqa: E731
        return JaggedTensor(values=mv(self._values), weights=mv(self._weights), lengths=mv(self._lengths),
                            offsets=mv(self._offsets))

    def __str__(self) -> str:
        return f"JaggedTensor(values={self._values}, lengths={self.lengths()}, weights={self._weights})"


class KeyedJaggedTensor(_Jagged):
    """Jagged tensor with a key (feature) dimension: F keys x B samples x variable length."""

    def __init__(self, keys: List[str], values: torch.Tensor, weights: Optional[torch.Tensor] = None,
                 lengths: Optional[torch.Tensor] = None, offsets: Optional[torch.Tensor] = None,
                 stride: Optional[int] = None, length_per_key: Optional[List[int]] = None,
                 offset_per_key: Optional[List[int]] = None, index_per_key: Optional[Dict[str, int]] = None,
                 jt_dict: Optional[Dict[str, JaggedTensor]] = None) -> None:
        self._keys: List[str] = keys
        self._init_jagged(values, weights, lengths, offsets)
        if stride is None:
            if len(keys) == 0:
                stride = 0
            elif offsets is not None and offsets.numel() > 0:
                stride = (offsets.numel() - 1) // len(keys)
            elif lengths is not None:
                stride = lengths.numel() // len(keys)
            else:
                stride = 0
        self._stride: int = stride
        self._length_per_key: Optional[List[int]] = length_per_key
        self._offset_per_key: Optional[List[int]] = offset_per_key
        self._index_per_key: Optional[Dict[str, int]] = index_per_key
        self._jt_dict: Optional[Dict[str, JaggedTensor]] = jt_dict

    # ---- constructors
    @staticmethod
    def from_offsets_sync(keys, values, offsets, weights=None, stride=None) -> "KeyedJaggedTensor":
        return KeyedJaggedTensor(keys=keys, values=values, weights=weights, offsets=offsets, stride=stride).sync()

    @staticmethod
    def from_lengths_sync(keys, values, lengths, weights=None, stride=None) -> "KeyedJaggedTensor":
        return KeyedJaggedTensor(keys=keys, values=values, weights=weights, lengths=lengths, stride=stride).sync()

    @staticmethod
    def concat(a: "KeyedJaggedTensor", b: "KeyedJaggedTensor") -> "KeyedJaggedTensor":
        if a.stride() != b.stride():
            raise ValueError("Can only merge KJTs of the same stride")
        w = None
        if a.weights_or_none() is not None and b.weights_or_none() is not None:
            w = torch.cat([a.weights(), b.weights()])
        return KeyedJaggedTensor(keys=a.keys() + b.keys(), values=torch.cat([a.values(), b.values()]), weights=w,
                                 lengths=torch.cat([a.lengths(), b.lengths()]), stride=a.stride())

    @staticmethod
    def empty(is_weighted: bool = False, device: Optional[torch.device] = None,
              values_dtype: Optional[torch.dtype] = None, weights_dtype: Optional[torch.dtype] = None,
              lengths_dtype: torch.dtype = torch.int32) -> "KeyedJaggedTensor":
        return KeyedJaggedTensor(
            keys=[], values=torch.tensor([], device=device, dtype=values_dtype),
            weights=torch.tensor([], device=device, dtype=weights_dtype) if is_weighted else None,
            lengths=torch.tensor([], device=device, dtype=lengths_dtype), stride=0)

    @staticmethod
    def empty_like(kjt: "KeyedJaggedTensor") -> "KeyedJaggedTensor":
        return KeyedJaggedTensor.empty(kjt.weights_or_none() is not None, kjt.device(), kjt.values().dtype,
                                       None if kjt.weights_or_none() is None else kjt.weights().dtype,
                                       kjt.lengths().dtype)

    # ---- accessors
    def sync(self) -> "KeyedJaggedTensor":
        self.length_per_key()
        self.offset_per_key()
        return self

    def device(self) -> torch.device:
        return self._values.device

    def keys(self) -> List[str]:
        return self._keys

    def stride(self) -> int:
        return self._stride

    def _key_indices(self) -> Dict[str, int]:
        if self._index_per_key is None:
            self._index_per_key = {k: i for i, k in enumerate(self._keys)}
        return self._index_per_key

    def length_per_key(self) -> List[int]:
        if self._length_per_key is None:
            if len(self._keys) == 0:
                self._length_per_key = []
            else:
                # the one host sync of the type (reference :502-508); avoided when the caller
                # passed length_per_key to the constructor
                self._length_per_key = self.lengths().view(len(self._keys), -1).sum(dim=1).tolist()
        return self._length_per_key

    def offset_per_key(self) -> List[int]:
        if self._offset_per_key is None:
            self._offset_per_key = _prefix(self.length_per_key())
        return self._offset_per_key

    # ---- structure ops
    def split(self, segments: List[int]) -> List["KeyedJaggedTensor"]:
        out: List[KeyedJaggedTensor] = []
        lpk, opk = self.length_per_key(), self.offset_per_key()
        k0 = 0
        for seg in segments:
            k1 = k0 + seg
            if seg == len(self._keys):
                out.append(KeyedJaggedTensor(self._keys, self._values, self._weights, self._lengths, self._offsets,
                                             self._stride, self._length_per_key, self._offset_per_key,
                                             self._index_per_key, self._jt_dict))
            elif seg == 0:
                e = KeyedJaggedTensor.empty_like(self)
                out.append(KeyedJaggedTensor(keys=[], values=e.values(), weights=e.weights_or_none(),
                                             lengths=e.lengths(), stride=self._stride))
            else:
                v0, v1 = opk[k0], opk[k1]
                out.append(KeyedJaggedTensor(
                    keys=self._keys[k0:k1], values=self._values[v0:v1],
                    weights=None if self._weights is None else self._weights[v0:v1],
                    lengths=self.lengths()[k0 * self._stride:k1 * self._stride], stride=self._stride,
                    length_per_key=lpk[k0:k1]))
            k0 = k1
        return out

    def permute(self, indices: List[int], indices_tensor: Optional[torch.Tensor] = None) -> "KeyedJaggedTensor":
        if indices_tensor is None:
            indices_tensor = torch.tensor(indices, dtype=torch.int32, device=self.device())
        lpk = self.length_per_key()
        new_keys = [self._keys[i] for i in indices]
        new_lpk = [lpk[i] for i in indices]
        nl, nv, nw = torch.ops.fbgemm.permute_2D_sparse_data(
            indices_tensor, self.lengths().view(len(self._keys), -1), self._values, self._weights, sum(new_lpk))
        return KeyedJaggedTensor(keys=new_keys, values=nv, weights=nw, lengths=nl.view(-1), stride=self._stride,
                                 length_per_key=new_lpk if new_keys else None)

    def __getitem__(self, key: str) -> JaggedTensor:
        i = self._key_indices()[key]
        opk = self.offset_per_key()
        return JaggedTensor(values=self._values[opk[i]:opk[i + 1]],
                            weights=None if self._weights is None else self._weights[opk[i]:opk[i + 1]],
                            lengths=self.lengths()[i * self._stride:(i + 1) * self._stride])

    def to_dict(self) -> Dict[str, JaggedTensor]:
        if self._jt_dict is None:
            opk = self.offset_per_key()
            lengths = self.lengths()
            d: Dict[str, JaggedTensor] = {}
            for i, k in enumerate(self._keys):
                seg_len = lengths[i * self._stride:(i + 1) * self._stride]
                d[k] = JaggedTensor(values=self._values[opk[i]:opk[i + 1]],
                                    weights=None if self._weights is None else self._weights[opk[i]:opk[i + 1]],
                                    lengths=seg_len, offsets=_offsets_of(seg_len))
            self._jt_dict = d
        return self._jt_dict

    def to(self, device: torch.device, non_blocking: bool = False) -> "KeyedJaggedTensor":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)  # noqa: E731
        return KeyedJaggedTensor(keys=self._keys, values=mv(self._values), weights=mv(self._weights),
                                 lengths=mv(self._lengths), offsets=mv(self._offsets), stride=self._stride,
                                 length_per_key=self._length_per_key, offset_per_key=self._offset_per_key,
                                 index_per_key=self._index_per_key)

    def pin_memory(self) -> "KeyedJaggedTensor":
        pin = lambda t: None if t is None else t.pin_memory()  # noqa: E731
        return KeyedJaggedTensor(keys=self._keys, values=pin(self._values), weights=pin(self._weights),
                                 lengths=pin(self._lengths), offsets=pin(self._offsets), stride=self._stride,
                                 length_per_key=self._length_per_key, offset_per_key=self._offset_per_key)

    def __str__(self) -> str:
        if len(self._keys) == 0:
            return "KeyedJaggedTensor()\n"
        return "KeyedJaggedTensor({\n" + "".join(f'    "{k}": {jt},\n' for k, jt in self.to_dict().items()) + "})\n"


class KeyedTensor:
    """Dense [B, sum_k D_k] tensor with named column blocks (reference :1101)."""

    def __init__(self, keys: List[str], length_per_key: List[int], values: torch.Tensor, key_dim: int = 1,
                 offset_per_key: Optional[List[int]] = None, index_per_key: Optional[Dict[str, int]] = None) -> None:
        self._keys, self._length_per_key, self._values, self._key_dim = keys, length_per_key, values, key_dim
        self._offset_per_key, self._index_per_key = offset_per_key, index_per_key

    @staticmethod
    def from_tensor_list(keys: List[str], tensors: List[torch.Tensor], key_dim: int = 1,
                         cat_dim: int = 1) -> "KeyedTensor":
        return KeyedTensor(keys=keys, length_per_key=[t.shape[key_dim] for t in tensors],
                           values=torch.cat(tensors, dim=cat_dim), key_dim=key_dim)

    def keys(self) -> List[str]:
        return self._keys

    def values(self) -> torch.Tensor:
        return self._values

    def key_dim(self) -> int:
        return self._key_dim

    def length_per_key(self) -> List[int]:
        return self._length_per_key

    def offset_per_key(self) -> List[int]:
        if self._offset_per_key is None:
            self._offset_per_key = _prefix(self._length_per_key)
        return self._offset_per_key

    def _key_indices(self) -> Dict[str, int]:
        if self._index_per_key is None:
            self._index_per_key = {k: i for i, k in enumerate(self._keys)}
        return self._index_per_key

    def __getitem__(self, key: str) -> torch.Tensor:
        i = self._key_indices()[key]
        o = self.offset_per_key()
        return self._values.narrow(self._key_dim, o[i], self._length_per_key[i])

    def to_dict(self) -> Dict[str, torch.Tensor]:
        parts = torch.split(self._values, self._length_per_key, dim=self._key_dim)
        return dict(zip(self._keys, parts))

    @staticmethod
    def regroup(keyed_tensors: List["KeyedTensor"], groups: List[List[str]]) -> List[torch.Tensor]:
        lookup: Dict[str, Tuple[int, int, int]] = {}
        key_dim = keyed_tensors[0].key_dim()
        for ti, kt in enumerate(keyed_tensors):
            o = kt.offset_per_key()
            for ki, k in enumerate(kt.keys()):
                lookup[k] = (ti, o[ki], kt.length_per_key()[ki])
        out = []
        for g in groups:
            cols = [keyed_tensors[lookup[k][0]].values().narrow(key_dim, lookup[k][1], lookup[k][2]) for k in g]
            out.append(torch.cat(cols, dim=key_dim))
        return out

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        self._values.record_stream(stream)

    def to(self, device: torch.device, non_blocking: bool = False) -> "KeyedTensor":
        return KeyedTensor(self._keys, self._length_per_key, self._values.to(device, non_blocking=non_blocking),
                           self._key_dim, self._offset_per_key, self._index_per_key)

    def __str__(self) -> str:
        return "KeyedTensor({\n" + "".join(f'    "{k}": {v},\n' for k, v in self.to_dict().items()) + "})\n"
