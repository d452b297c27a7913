"""pytest glue replacing ``pytest-mpi`` (reference requirements.txt:9; README.md:189,200,212):

* ``--with-mpi`` flag and ``@pytest.mark.mpi``: marked tests are skipped unless the flag is
  given, exactly the plugin behaviour the reference's test commands rely on;
* under the launcher every rank runs its own pytest; ranks other than 0 keep quiet unless
  something fails.
"""
import os

import pytest


def pytest_addoption(parser):
    group = parser.getgroup("mpi", "ccmpi_b200 multi-rank testing")
    try:
        group.addoption("--with-mpi", action="store_true", default=False,
                        help="run tests marked 'mpi' (start pytest under the launcher: "
                             "python -m collective_communication_mpi_b200.launch -n N python -m pytest --with-mpi ...)")
    except ValueError:
        pass  # a real pytest-mpi is installed and already owns the flag


def pytest_configure(config):
    config.addinivalue_line("markers", "mpi: test needs several ranks; runs only with --with-mpi")
    config.addinivalue_line("markers", "mpi_skip: skip when running with --with-mpi")
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200)")
    if int(os.environ.get("RANK", "0")) != 0 and os.environ.get("CCMPI_TEST_VERBOSE_ALL") != "1":
        # one readable report per job: non-root ranks print only failures
        config.option.verbose = min(config.option.verbose, 0)
        rep = config.pluginmanager.getplugin("terminalreporter")
        if rep is not None:
            config.option.quiet = True if hasattr(config.option, "quiet") else None


def pytest_collection_modifyitems(config, items):
    with_mpi = config.getoption("--with-mpi", default=False)
    skip_mpi = pytest.mark.skip(reason="needs --with-mpi and a multi-rank launch")
    skip_single = pytest.mark.skip(reason="skipped under --with-mpi")
    for item in items:
        if "mpi" in item.keywords and not with_mpi:
            item.add_marker(skip_mpi)
        if "mpi_skip" in item.keywords and with_mpi:
            item.add_marker(skip_single)
