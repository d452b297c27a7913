"""DP sharding (reference behaviour: data/data_parallel_preprocess.py:3-59; its tests use the
grids (mp,dp) = (2,1), (1,2), (2,2), (2,4) on 8 samples -- tests/test_data_split.py:71-200)."""
import numpy as np
import pytest

from data.data_parallel_preprocess import split_data

N = 8
X = np.arange(N * 2 * 2, dtype=np.float32).reshape(N, 2, 2)
Y = np.arange(N * 2, dtype=np.int32).reshape(N, 2)


@pytest.mark.parametrize("mp_size,dp_size", [(2, 1), (1, 2), (2, 2), (2, 4), (4, 2), (1, 8), (8, 1)])
def test_every_rank_gets_its_replica_slice(mp_size, dp_size):
    per = N // dp_size
    for rank in range(mp_size * dp_size):
        xs, ys = split_data(x_train=X, y_train=Y, mp_size=mp_size, dp_size=dp_size, rank=rank)
        replica = rank // mp_size
        assert xs.shape[0] * dp_size == N and ys.shape[0] * dp_size == N
        np.testing.assert_allclose(xs, X[replica * per:(replica + 1) * per])
        np.testing.assert_allclose(ys, Y[replica * per:(replica + 1) * per])


def test_mp_ranks_of_a_replica_share_data_and_nothing_is_copied():
    a, _ = split_data(X, Y, mp_size=2, dp_size=4, rank=2)
    b, _ = split_data(X, Y, mp_size=2, dp_size=4, rank=3)
    np.testing.assert_array_equal(a, b)
    assert np.shares_memory(a, X), "shards are views of the dataset, like the reference's slices"


def test_order_is_preserved_and_shards_tile_the_dataset():
    parts = [split_data(X, Y, mp_size=1, dp_size=4, rank=r)[0] for r in range(4)]
    np.testing.assert_array_equal(np.concatenate(parts), X)


def test_torch_tensors_are_sliced_the_same_way():
    torch = pytest.importorskip("torch")
    xs, ys = split_data(torch.from_numpy(X), torch.from_numpy(Y), mp_size=2, dp_size=2, rank=3)
    np.testing.assert_array_equal(xs.numpy(), X[4:8])
    np.testing.assert_array_equal(ys.numpy(), Y[4:8])
