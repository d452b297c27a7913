"""Reference-compatible ``split_data`` (reference data/data_parallel_preprocess.py:3); the
implementation is ``collective_communication_mpi_b200.parallel.layout.split_data``. Like the
reference, ``data`` is a namespace package (no ``__init__.py``)."""
from collective_communication_mpi_b200.parallel import layout as _layout


def split_data(x_train, y_train, mp_size: int, dp_size: int, rank: int):
    return _layout.split_data(x_train, y_train, mp_size=mp_size, dp_size=dp_size, rank=rank)


split_data.__doc__ = _layout.split_data.__doc__
