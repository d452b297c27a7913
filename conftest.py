"""Root conftest: loads the multi-rank pytest glue (``--with-mpi`` / ``@pytest.mark.mpi``) so the
reference's README test commands work with the launcher in place of ``mpirun``."""
from collective_communication_mpi_b200.pytest_plugin import (  # noqa: F401
    pytest_addoption, pytest_collection_modifyitems, pytest_configure,
)
