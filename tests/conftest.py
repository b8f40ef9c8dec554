import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver via gpurun)")
    config.addinivalue_line("markers", "slow: long-running CPU integration test")


def pytest_collection_modifyitems(config, items):
    # tests/test_zz_* files are collected last by name; keep that order explicit.
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz"),))


@pytest.fixture(scope="session")
def ssd_files(tmp_path_factory):
    """Prototxt + random-init caffemodel (conf gain 2.5 => plenty of detections)."""
    from distributedvolunteercomputing_b200 import caffe_io
    d = tmp_path_factory.mktemp("ssd")
    proto = d / "MobileNetSSD_deploy.prototxt.txt"
    model = d / "MobileNetSSD_deploy.caffemodel"
    proto.write_text(caffe_io.mobilenet_ssd_prototxt())
    net = caffe_io.load_netdef(str(proto))
    w = caffe_io.random_init_weights(net, seed=1, conf_gain=2.5)
    caffe_io.write_caffemodel(str(model), w)
    return str(proto), str(model)


@pytest.fixture(scope="session")
def graph(ssd_files):
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
    return SSDGraph.from_files(*ssd_files)
