"""tf.data-like input pipeline with global-batch semantics under a strategy.

Reference call sites: ``map(scale).cache().shuffle(10000)`` then
``.batch(GLOBAL_BATCH_SIZE)`` + ``Options().experimental_distribute.auto_shard_policy = OFF`` +
``with_options`` (distributed_with_keras.py:18-30,54-57); ``from_tensor_slices`` ->
``shuffle(1000).repeat().batch(bs).prefetch(100)`` for TRAIN, ``batch(bs)`` for EVAL/PREDICT
(mnist_keras_distributed.py:123-148).

Design: datasets built from in-memory tensors stay *index based* as long as possible -- shuffle,
repeat, take, skip and shard transform a stream of row indices, and ``batch`` gathers whole batches
with one vectorised copy (numpy) or, when the native runtime is built, in a C++ worker thread that
fills a ring of pinned host buffers (``csrc/runtime/pipeline.cc``) from which the H2D copy is issued
directly. Anything else (``map`` after ``batch``, python generators) falls back to a generic
element-at-a-time iterator with identical semantics.

Shuffle follows tf.data: a buffer of ``buffer_size`` elements is filled from upstream, each output is
drawn uniformly from the buffer and replaced by the next upstream element; reshuffled every epoch.
"""
from __future__ import annotations

import queue
import threading
from typing import Callable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch


class AutoShardPolicy:
    OFF = "OFF"
    AUTO = "AUTO"
    FILE = "FILE"
    DATA = "DATA"


class _DistributeOptions:
    def __init__(self):
        self.auto_shard_policy = AutoShardPolicy.AUTO


class Options:
    """tf.data.Options with the one knob the reference sets (distributed_with_keras.py:55-56)."""

    def __init__(self):
        self.experimental_distribute = _DistributeOptions()
        self.experimental_deterministic = True


def _to_numpy(a):
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


def _stack(elems):
    first = elems[0]
    if isinstance(first, tuple):
        return tuple(np.stack([e[i] for e in elems]) for i in range(len(first)))
    return np.stack(elems)


class _Source:
    """Random-access element store: a tuple of equally long arrays (or a single array)."""

    def __init__(self, arrays, is_tuple):
        self.arrays = [np.ascontiguousarray(a) for a in arrays]
        self.is_tuple = is_tuple
        self.n = len(self.arrays[0])
        for a in self.arrays:
            if len(a) != self.n:
                raise ValueError("from_tensor_slices: all tensors must have the same first dimension")

    def get(self, i):
        out = tuple(a[i] for a in self.arrays)
        return out if self.is_tuple else out[0]

    def gather(self, idx):
        out = tuple(a[idx] for a in self.arrays)
        return out if self.is_tuple else out[0]


class _MappedSource:
    def __init__(self, src, fn):
        self.src, self.fn, self.n = src, fn, src.n
        self.is_tuple = True

    def get(self, i):
        e = self.src.get(i)
        r = self.fn(*e) if isinstance(e, tuple) else self.fn(e)
        return tuple(_to_numpy(x) for x in r) if isinstance(r, tuple) else _to_numpy(r)

    def gather(self, idx):
        return _stack([self.get(int(i)) for i in idx])

    def materialize(self) -> _Source:
        elems = [self.get(i) for i in range(self.n)]
        st = _stack(elems)
        return _Source(list(st) if isinstance(st, tuple) else [st], isinstance(st, tuple))


class Dataset:
    """Lazily evaluated pipeline node. Iterate it to get numpy elements (tuples for structured data)."""

    def __init__(self, kind: str, parent: Optional["Dataset"] = None, **p):
        self.kind, self.parent, self.p = kind, parent, p
        self._options: Optional[Options] = parent._options if parent is not None else None
        self._cache_store = None
        self._epoch = 0

    # ---- constructors ------------------------------------------------------------------------
    @staticmethod
    def from_tensor_slices(tensors) -> "Dataset":
        if isinstance(tensors, (tuple, list)):
            return Dataset("source", src=_Source([_to_numpy(t) for t in tensors], True))
        return Dataset("source", src=_Source([_to_numpy(tensors)], False))

    @staticmethod
    def from_generator(gen_fn: Callable[[], Iterator]) -> "Dataset":
        return Dataset("generator", gen=gen_fn)

    @staticmethod
    def range(n: int) -> "Dataset":
        return Dataset.from_tensor_slices(np.arange(n, dtype=np.int64))

    # ---- transformations ---------------------------------------------------------------------
    def map(self, fn, num_parallel_calls=None) -> "Dataset":
        return Dataset("map", self, fn=fn)

    def cache(self, filename: str = "") -> "Dataset":
        return Dataset("cache", self)

    def shuffle(self, buffer_size: int, seed: Optional[int] = None, reshuffle_each_iteration=True) -> "Dataset":
        return Dataset("shuffle", self, buffer=int(buffer_size), seed=seed, reshuffle=reshuffle_each_iteration)

    def repeat(self, count: Optional[int] = None) -> "Dataset":
        return Dataset("repeat", self, count=count)

    def batch(self, batch_size: int, drop_remainder: bool = False) -> "Dataset":
        return Dataset("batch", self, batch=int(batch_size), drop=bool(drop_remainder))

    def prefetch(self, buffer_size: int) -> "Dataset":
        return Dataset("prefetch", self, depth=max(1, int(buffer_size)) if buffer_size and buffer_size > 0 else 2)

    def take(self, count: int) -> "Dataset":
        return Dataset("take", self, count=int(count))

    def skip(self, count: int) -> "Dataset":
        return Dataset("skip", self, count=int(count))

    def shard(self, num_shards: int, index: int) -> "Dataset":
        if not 0 <= index < num_shards:
            raise ValueError("shard index out of range")
        return Dataset("shard", self, num=int(num_shards), index=int(index))

    def with_options(self, options: Options) -> "Dataset":
        d = Dataset("options", self)
        d._options = options
        return d

    def options(self) -> Options:
        return self._options if self._options is not None else Options()

    # ---- introspection ------------------------------------------------------------------------
    def _chain(self) -> List["Dataset"]:
        out, d = [], self
        while d is not None:
            out.append(d)
            d = d.parent
        return out[::-1]

    @property
    def global_batch_size(self) -> Optional[int]:
        for d in self._chain()[::-1]:
            if d.kind == "batch":
                return d.p["batch"]
        return None

    def cardinality(self) -> int:
        """Number of elements, -1 if infinite, -2 if unknown."""
        k, p = self.kind, self.p
        if k == "source":
            return p["src"].n
        if k == "generator":
            return -2
        c = self.parent.cardinality()
        if k in ("map", "cache", "shuffle", "prefetch", "options"):
            return c
        if k == "repeat":
            return -1 if p["count"] is None else (c * p["count"] if c >= 0 else c)
        if c < 0:
            return c
        if k == "batch":
            return c // p["batch"] if p["drop"] else -(-c // p["batch"])
        if k == "take":
            return min(c, p["count"])
        if k == "skip":
            return max(c - p["count"], 0)
        if k == "shard":
            return (c - p["index"] + p["num"] - 1) // p["num"]
        return -2

    def __len__(self):
        c = self.cardinality()
        if c < 0:
            raise TypeError("dataset is infinite or of unknown length")
        return c

    # ---- index-based fast path ------------------------------------------------------------------
    def _index_plan(self):
        """Returns (source, index_iterator_factory) when this node is expressible on row indices."""
        k, p = self.kind, self.p
        if k == "source":
            src = p["src"]
            return src, (lambda: iter(range(src.n)))
        if self.parent is None:
            return None
        up = self.parent._index_plan()
        if up is None:
            return None
        src, mk = up
        if k == "options":
            return src, mk
        if k == "map":
            return _MappedSource(src, p["fn"]), mk
        if k == "cache":
            if self._cache_store is None:
                # cache() right after from_tensor_slices/map: materialise once, keep the index stream
                self._cache_store = src.materialize() if isinstance(src, _MappedSource) else src
            return self._cache_store, mk
        if k == "shuffle":
            node = self

            def mk_shuffle():
                seed = p["seed"]
                epoch = node._epoch
                node._epoch += 1
                if seed is None:
                    rng = np.random.default_rng()
                else:
                    rng = np.random.default_rng([seed, epoch if p["reshuffle"] else 0])
                return _shuffle_iter(mk(), p["buffer"], rng)
            return src, mk_shuffle
        if k == "repeat":
            def mk_repeat():
                cnt = p["count"]
                i = 0
                while cnt is None or i < cnt:
                    empty = True
                    for v in mk():
                        empty = False
                        yield v
                    if empty:
                        return
                    i += 1
            return src, mk_repeat
        if k == "take":
            def mk_take():
                for j, v in enumerate(mk()):
                    if j >= p["count"]:
                        return
                    yield v
            return src, mk_take
        if k == "skip":
            def mk_skip():
                for j, v in enumerate(mk()):
                    if j >= p["count"]:
                        yield v
            return src, mk_skip
        if k == "shard":
            def mk_shard():
                for j, v in enumerate(mk()):
                    if j % p["num"] == p["index"]:
                        yield v
            return src, mk_shard
        return None

    # ---- native (C++) fast path -------------------------------------------------------------------
    def _native_iter(self):
        """``source [-> shard] [-> shuffle] [-> repeat] -> batch [-> prefetch]`` over plain arrays runs in
        the C++ runtime (background thread + prefetch ring). Returns None when the chain does not match or
        the runtime library is not built."""
        from . import native_pipeline as npipe
        if not npipe.available():
            return None
        chain = [d for d in self._chain() if d.kind not in ("options", "cache")]
        depth = 2
        if chain and chain[-1].kind == "prefetch":
            depth = min(max(2, chain[-1].p["depth"]), 64)
            chain = chain[:-1]
        kinds = [d.kind for d in chain]
        if len(kinds) < 2 or kinds[0] != "source" or kinds[-1] != "batch":
            return None
        mid = kinds[1:-1]
        allowed = ["shard", "shuffle", "repeat"]
        if any(k not in allowed for k in mid) or [k for k in allowed if k in mid] != mid:
            return None
        src = chain[0].p["src"]
        if len(src.arrays) > 2:
            return None
        node = {d.kind: d for d in chain}
        shuffle = node.get("shuffle")
        repeat = node.get("repeat")
        shard = node.get("shard")
        seed = None
        if shuffle is not None and shuffle.p["seed"] is not None:
            seed = (int(shuffle.p["seed"]) * 1000003 + (shuffle._epoch if shuffle.p["reshuffle"] else 0)) & (2 ** 63 - 1)
            shuffle._epoch += 1
        rep = 1 if repeat is None else (-1 if repeat.p["count"] is None else int(repeat.p["count"]))
        try:
            return npipe.NativeBatchIterator(
                src.arrays[0], src.arrays[1] if len(src.arrays) == 2 else None, node["batch"].p["batch"],
                shuffle_buffer=0 if shuffle is None else shuffle.p["buffer"], seed=seed, repeat=rep,
                drop_remainder=node["batch"].p["drop"], depth=depth,
                shard_num=1 if shard is None else shard.p["num"], shard_index=0 if shard is None else shard.p["index"])
        except (TypeError, OSError, RuntimeError):
            return None

    # ---- iteration ----------------------------------------------------------------------------
    def __iter__(self):
        if self.kind in ("batch", "prefetch", "options"):
            it = self._native_iter()
            if it is not None:
                return it
        return self._iter()

    def _iter(self):
        k, p = self.kind, self.p
        plan = self._index_plan() if k not in ("batch", "prefetch", "generator") else None
        if plan is not None:
            src, mk = plan
            for i in mk():
                yield src.get(int(i))
            return
        if k == "generator":
            yield from p["gen"]()
        elif k == "batch":
            up = self.parent._index_plan()
            bs, drop = p["batch"], p["drop"]
            if up is not None:
                src, mk = up
                buf = []
                for i in mk():
                    buf.append(i)
                    if len(buf) == bs:
                        yield src.gather(np.asarray(buf, dtype=np.int64))
                        buf = []
                if buf and not drop:
                    yield src.gather(np.asarray(buf, dtype=np.int64))
            else:
                buf = []
                for e in self.parent._iter():
                    buf.append(e)
                    if len(buf) == bs:
                        yield _stack(buf)
                        buf = []
                if buf and not drop:
                    yield _stack(buf)
        elif k == "prefetch":
            yield from _prefetch_iter(self.parent._iter, p["depth"])
        elif k == "map":
            for e in self.parent._iter():
                r = p["fn"](*e) if isinstance(e, tuple) else p["fn"](e)
                yield tuple(_to_numpy(x) for x in r) if isinstance(r, tuple) else _to_numpy(r)
        elif k == "cache":
            if self._cache_store is None:
                store = []
                for e in self.parent._iter():
                    store.append(e)
                    yield e
                self._cache_store = store
            else:
                yield from self._cache_store
        elif k == "shuffle":
            seed = p["seed"]
            epoch = self._epoch
            self._epoch += 1
            rng = np.random.default_rng() if seed is None else np.random.default_rng(
                [seed, epoch if p["reshuffle"] else 0])
            yield from _shuffle_iter(self.parent._iter(), p["buffer"], rng)
        elif k == "repeat":
            cnt, i = p["count"], 0
            while cnt is None or i < cnt:
                empty = True
                for e in self.parent._iter():
                    empty = False
                    yield e
                if empty:
                    return
                i += 1
        elif k == "take":
            for j, e in enumerate(self.parent._iter()):
                if j >= p["count"]:
                    return
                yield e
        elif k == "skip":
            for j, e in enumerate(self.parent._iter()):
                if j >= p["count"]:
                    yield e
        elif k == "shard":
            for j, e in enumerate(self.parent._iter()):
                if j % p["num"] == p["index"]:
                    yield e
        elif k == "options":
            yield from self.parent._iter()
        else:
            raise RuntimeError(f"unknown dataset node {k}")

    def as_numpy_iterator(self):
        return iter(self)


def _shuffle_iter(upstream, buffer_size: int, rng):
    buf = []
    for v in upstream:
        if len(buf) < buffer_size:
            buf.append(v)
            continue
        j = int(rng.integers(0, len(buf)))
        out, buf[j] = buf[j], v
        yield out
    while buf:
        j = int(rng.integers(0, len(buf)))
        buf[j], buf[-1] = buf[-1], buf[j]
        yield buf.pop()


def _prefetch_iter(make_iter, depth: int):
    q: "queue.Queue" = queue.Queue(maxsize=depth)
    stop = threading.Event()
    END = object()

    def worker():
        try:
            for e in make_iter():
                while not stop.is_set():
                    try:
                        q.put(e, timeout=0.1)
                        break
                    except queue.Full:
                        continue
                if stop.is_set():
                    return
            q.put(END)
        except BaseException as ex:  # surface producer errors in the consumer
            q.put(ex)

    t = threading.Thread(target=worker, daemon=True, name="tfde-prefetch")
    t.start()
    try:
        while True:
            e = q.get()
            if e is END:
                return
            if isinstance(e, BaseException):
                raise e
            yield e
    finally:
        stop.set()


# ----------------------------------------------------------------------------------------------
# distribution: global batch -> per-replica batches
# ----------------------------------------------------------------------------------------------
class DistributedDataset:
    """What ``strategy.experimental_distribute_dataset`` / ``fit`` iterate: yields
    ``(x_replica, y_replica, global_batch)`` for THIS replica.

    The dataset is batched with the GLOBAL batch size (distributed_with_keras.py:15,54); every global
    batch is re-batched into ``num_replicas`` equal per-replica batches (TF's rebatching).

    * ``AutoShardPolicy.OFF`` (distributed_with_keras.py:55-57): no sharding -- every worker iterates
      the whole dataset in its own (unseeded => different) order and consumes *all* sub-batches of its
      own stream, one per step.
    * ``DATA`` (and ``AUTO``/``FILE`` for in-memory data, which has no files to shard): sub-batch ``r`` of
      every global batch goes to replica ``r``; disjoint coverage needs the same element order on every
      worker, i.e. a seeded shuffle.
    * evaluation (``for_eval``): always DATA-style slicing of each batch, including the short last
      batch (mnist_keras_distributed.py:146-147 keeps the remainder), so every example is scored once.
    """

    def __init__(self, dataset: Dataset, strategy, for_eval: bool = False):
        self.dataset = dataset
        self.strategy = strategy
        self.for_eval = for_eval
        pol = dataset.options().experimental_distribute.auto_shard_policy
        self.policy = AutoShardPolicy.DATA if pol in (AutoShardPolicy.AUTO, AutoShardPolicy.FILE) else pol
        if for_eval:
            self.policy = AutoShardPolicy.DATA

    def __iter__(self):
        n = self.strategy.num_replicas_in_sync
        r = self.strategy.replica_id
        for elem in self.dataset:
            if not isinstance(elem, tuple) or len(elem) < 2:
                raise ValueError("training datasets must yield (features, labels)")
            x, y = elem[0], elem[1]
            gb = int(x.shape[0])
            if n == 1:
                yield _as_tensor(x), _as_tensor(y), gb
                continue
            if self.for_eval:
                per = -(-gb // n)
                lo, hi = min(r * per, gb), min((r + 1) * per, gb)
                yield _as_tensor(x[lo:hi]), _as_tensor(y[lo:hi]), gb
                continue
            per = gb // n
            if per == 0:
                continue
            if per * n != gb:
                gb = per * n   # drop the ragged tail so replicas stay in lock-step
            if self.policy == AutoShardPolicy.OFF:
                for s in range(n):
                    yield _as_tensor(x[s * per:(s + 1) * per]), _as_tensor(y[s * per:(s + 1) * per]), gb
            else:
                yield _as_tensor(x[r * per:(r + 1) * per]), _as_tensor(y[r * per:(r + 1) * per]), gb


def _as_tensor(a):
    if isinstance(a, torch.Tensor):
        return a
    a = np.ascontiguousarray(a)
    if a.dtype == np.float64:
        a = a.astype(np.float32)   # the reference feeds float64 numpy (images/255.0); TF casts to float32
    return torch.from_numpy(a)


def as_distributed(ds, strategy, for_eval=False) -> DistributedDataset:
    if isinstance(ds, DistributedDataset):
        return ds
    return DistributedDataset(ds, strategy, for_eval=for_eval)
