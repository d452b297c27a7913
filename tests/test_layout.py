"""Grid arithmetic of ``get_info`` that needs no communication (reference func_impl.py:53-72)."""
import pytest

from collective_communication_mpi_b200.parallel.layout import GridPos, partition_dims


@pytest.mark.parametrize("layer", ["fc_q", "fc_k", "fc_v", "fc1"])
def test_column_parallel_layers_split_the_output(layer):
    assert partition_dims(layer, 768, 256, 4) == (768, 64)


@pytest.mark.parametrize("layer", ["fc_o", "fc2"])
def test_row_parallel_layers_split_the_input(layer):
    assert partition_dims(layer, 256, 10, 4) == (64, 10)


def test_unknown_layer_is_rejected_like_the_reference():
    with pytest.raises(ValueError, match="Invalid fc_layer: fc_z."):
        partition_dims("fc_z", 8, 8, 2)


def test_rank_map_is_model_parallel_major():
    # figs/group.png: mp_size=4, dp_size=2 -> ranks 0-3 are replica 0, 4-7 replica 1
    assert [GridPos(r, 4, 2).mp_idx for r in range(8)] == [0, 1, 2, 3, 0, 1, 2, 3]
    assert [GridPos(r, 4, 2).dp_idx for r in range(8)] == [0, 0, 0, 0, 1, 1, 1, 1]


def test_single_rank_get_info_needs_no_launcher():
    from mpi4py import MPI
    from model.func_impl import get_info
    mp_idx, dp_idx, mp_comm, dp_comm, pin, pout = get_info(comm=MPI.COMM_WORLD, rank=0, mp_size=1, dp_size=1,
                                                           fc_layer="fc_q", in_dim=12, out_dim=6)
    assert (mp_idx, dp_idx, pin, pout) == (0, 0, 12, 6)
    assert mp_comm.Get_size() == 1 and dp_comm.Get_size() == 1
