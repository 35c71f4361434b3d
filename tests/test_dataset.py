"""tf.data-like pipeline semantics."""
import numpy as np
import pytest

import tfde_b200 as tf
from tfde_b200.data.dataset import AutoShardPolicy, Dataset, DistributedDataset, Options


class FakeStrategy:
    def __init__(self, n, r):
        self.num_replicas_in_sync, self.replica_id = n, r


def _ds(n=10):
    return Dataset.from_tensor_slices((np.arange(n, dtype=np.float32).reshape(n, 1), np.arange(n, dtype=np.int64)))


def test_from_tensor_slices_and_batch_keeps_remainder():
    batches = list(_ds(10).batch(4))
    assert [len(b[0]) for b in batches] == [4, 4, 2]
    assert [len(b[0]) for b in _ds(10).batch(4, drop_remainder=True)] == [4, 4]
    assert len(_ds(10).batch(4)) == 3


def test_features_only_dataset():
    ds = Dataset.from_tensor_slices(np.arange(6)).batch(4)
    out = list(ds)
    assert out[0].tolist() == [0, 1, 2, 3] and out[1].tolist() == [4, 5]


def test_map_cache_runs_fn_once_per_element():
    calls = []

    def fn(x, y):
        calls.append(int(y))
        return x * 2, y

    ds = _ds(6).map(fn).cache()
    a = [int(e[1]) for e in ds]
    b = [float(e[0][0]) for e in ds]
    assert a == list(range(6)) and b == [0, 2, 4, 6, 8, 10]
    assert sorted(calls) == list(range(6)), "cache() must not re-run the map on the second pass"


def test_shuffle_is_a_permutation_and_window_bounded():
    n, buf = 100, 10
    out = [int(e[1]) for e in _ds(n).shuffle(buf, seed=1)]
    assert sorted(out) == list(range(n)) and out != list(range(n))
    # tf.data window semantics: element i cannot appear before position i - buffer
    for pos, v in enumerate(out):
        assert v <= pos + buf
    again = [int(e[1]) for e in _ds(n).shuffle(buf, seed=1)]
    assert again == out, "same seed, first epoch -> same order"


def test_shuffle_reshuffles_each_iteration():
    ds = _ds(50).shuffle(50, seed=3)
    assert [int(e[1]) for e in ds] != [int(e[1]) for e in ds]


def test_repeat_take_skip_shard_prefetch():
    assert [int(e[1]) for e in _ds(3).repeat(2)] == [0, 1, 2, 0, 1, 2]
    assert [int(e[1]) for e in _ds(3).repeat().take(7)] == [0, 1, 2, 0, 1, 2, 0]
    assert [int(e[1]) for e in _ds(5).skip(2)] == [2, 3, 4]
    assert [int(e[1]) for e in _ds(7).shard(3, 1)] == [1, 4]
    assert [b[1].tolist() for b in _ds(5).batch(2).prefetch(100)] == [[0, 1], [2, 3], [4]]
    assert _ds(3).repeat().cardinality() == -1


def test_train_input_fn_pipeline_of_the_reference():
    x = np.random.rand(50, 28, 28)
    y = np.arange(50).reshape(-1, 1)
    ds = Dataset.from_tensor_slices((x, y)).shuffle(1000).repeat().batch(16).prefetch(100)
    it = iter(ds)
    for _ in range(5):
        xb, yb = next(it)
        assert xb.shape == (16, 28, 28) and yb.shape == (16, 1)


def test_options_and_policy_default():
    ds = _ds(8).batch(4)
    assert ds.options().experimental_distribute.auto_shard_policy == AutoShardPolicy.AUTO
    o = Options()
    o.experimental_distribute.auto_shard_policy = tf.data.experimental.AutoShardPolicy.OFF
    assert ds.with_options(o).options().experimental_distribute.auto_shard_policy == "OFF"
    assert ds.with_options(o).global_batch_size == 4


def test_distributed_off_policy_rebatches_own_stream():
    o = Options()
    o.experimental_distribute.auto_shard_policy = AutoShardPolicy.OFF
    ds = _ds(8).batch(4).with_options(o)
    for r in range(2):
        got = [(yb.tolist(), gb) for _, yb, gb in DistributedDataset(ds, FakeStrategy(2, r))]
        assert got == [([0, 1], 4), ([2, 3], 4), ([4, 5], 4), ([6, 7], 4)]


def test_distributed_data_policy_is_disjoint_cover():
    o = Options()
    o.experimental_distribute.auto_shard_policy = AutoShardPolicy.DATA
    ds = _ds(8).batch(4).with_options(o)
    seen = []
    for r in range(2):
        for _, yb, gb in DistributedDataset(ds, FakeStrategy(2, r)):
            assert gb == 4 and len(yb) == 2
            seen += yb.tolist()
    assert sorted(seen) == list(range(8))


def test_distributed_eval_scores_every_example_once_including_short_batch():
    ds = _ds(10).batch(4)
    seen = []
    for r in range(3):
        for _, yb, _ in DistributedDataset(ds, FakeStrategy(3, r), for_eval=True):
            seen += yb.tolist()
    assert sorted(seen) == list(range(10))


def test_mnist_sources_shapes():
    (xtr, ytr), (xte, yte) = tf.keras.datasets.mnist.load_data(n_train=256, n_test=64)
    assert xtr.shape == (256, 28, 28) and xtr.dtype == np.uint8 and ytr.shape == (256,)
    dsets, info = tf.data.mnist.load("mnist", with_info=True, as_supervised=True, n_train=256, n_test=64)
    img, lab = next(iter(dsets["train"]))
    assert img.shape == (28, 28, 1) and info.splits["train"] == 256


def test_native_pipeline_matches_python_semantics():
    from tfde_b200.data import native_pipeline as npipe
    if not npipe.available():
        pytest.skip("runtime library not built")
    x = np.arange(103 * 3, dtype=np.float32).reshape(103, 3)
    y = np.arange(103, dtype=np.int64)
    ds = Dataset.from_tensor_slices((x, y)).shuffle(20, seed=5).repeat(2).batch(10).prefetch(4)
    it = iter(ds)
    assert type(it).__name__ == "NativeBatchIterator"
    batches = list(it)
    assert [len(b[1]) for b in batches] == [10] * 20 + [6]
    labels = np.concatenate([np.asarray(b[1]) for b in batches])
    feats = np.concatenate([np.asarray(b[0]) for b in batches])
    assert sorted(labels[:103].tolist()) == list(range(103)) and sorted(labels[103:].tolist()) == list(range(103))
    assert np.array_equal(feats[:, 0], labels * 3), "rows of every component stay aligned"
    for pos, v in enumerate(labels[:103]):
        assert v <= pos + 20, "windowed shuffle: element i cannot appear before position i - buffer"
    assert labels[:103].tolist() != labels[103:].tolist(), "reshuffled each epoch"
    # same seed -> same first-epoch order on a fresh dataset; shard + drop_remainder
    ds2 = Dataset.from_tensor_slices((x, y)).shuffle(20, seed=5).repeat(2).batch(10).prefetch(4)
    assert np.concatenate([np.asarray(b[1]) for b in ds2])[:103].tolist() == labels[:103].tolist()
    ds3 = Dataset.from_tensor_slices((x, y)).shard(4, 1).batch(8, drop_remainder=True)
    got = [np.asarray(b[1]).tolist() for b in ds3]
    assert got == [list(range(1, 32, 4)), list(range(33, 64, 4)), list(range(65, 96, 4))]
    # infinite repeat terminates cleanly when the consumer stops early
    it4 = iter(Dataset.from_tensor_slices((x, y)).shuffle(1000).repeat().batch(16).prefetch(100))
    for _ in range(30):
        xb, yb = next(it4)
        assert xb.shape == (16, 3)
    it4.close()


def test_python_fallback_when_native_disabled(monkeypatch):
    monkeypatch.setenv("TFDE_NATIVE_PIPELINE", "0")
    ds = _ds(10).shuffle(4, seed=1).batch(4)
    assert type(iter(ds)).__name__ == "generator"
    assert [len(b[0]) for b in ds] == [4, 4, 2]
