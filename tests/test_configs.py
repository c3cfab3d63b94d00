"""curvecloudnet_b200.configs rebuilds the shipped YAML `model:` sections; compare with the
reference's files whenever they are reachable (build container only)."""
import os

import pytest

from curvecloudnet_b200 import configs as C

REF = "/root/reference/configs/curvecloudnet-eval"
FILES = {"nuscenes": "nuscenes", "kitti": "kitti", "shapenet-class": "shapenet-class",
         "shapenet-seg": "shapenet-seg", "kortx": "kortx-testsplit", "a2d2": "audi"}


@pytest.mark.parametrize("name", list(FILES))
def test_tables_are_consistent(name):
    cfg = C.MODEL_CONFIGS[name][0]()
    n = len(cfg["steps"])
    for key in ("feat_dims", "knn", "ratios", "radii", "num_skips", "kernel_sizes"):
        assert len(cfg[key]) == n


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (GPU box)")
@pytest.mark.parametrize("name", list(FILES))
def test_tables_equal_reference_yaml(name):
    import yaml
    ref = yaml.safe_load(open(os.path.join(REF, FILES[name] + "-curvecloudnet.yaml")))["model"]
    mine = C.MODEL_CONFIGS[name][0]()
    assert set(ref) == set(mine)
    for k in ref:
        assert ref[k] == mine[k], k
