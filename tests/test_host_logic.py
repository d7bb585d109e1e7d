"""Host-side logic that needs no GPU: checkpoint reading; the resident map refuses CPU tensors."""
import numpy as np
import pytest
import torch

from fisher_nerf_customized_b200 import synth
from fisher_nerf_customized_b200.scoring import GaussianMap, read_npz_params


def test_read_npz_params_reads_the_reference_checkpoint_format(tmp_path):
    """Key set of save_params_ckpt (models/SLAM/utils/common_utils.py:45-59) incl. extras that must be ignored."""
    raw = synth.make_scene(2000, seed=5)
    path = tmp_path / "params7.npz"
    np.savez(path, **raw, cam_unnorm_rots=np.zeros((1, 4, 3), np.float32), cam_trans=np.zeros((1, 3, 3), np.float32),
             timestep=np.zeros(2000, np.float32), gt_w2c_all_frames=np.eye(4)[None])
    p = read_npz_params(str(path), device="cpu")
    assert sorted(p) == ["log_scales", "logit_opacities", "means3D", "rgb_colors", "unnorm_rotations"]
    for k, v in p.items():
        assert v.dtype == torch.float32
        np.testing.assert_array_equal(v.numpy(), np.asarray(raw[k], np.float32))
    bad = tmp_path / "x.npz"
    np.savez(bad, means3D=raw["means3D"])
    with pytest.raises(KeyError):
        read_npz_params(str(bad), device="cpu")


def test_gaussian_map_refuses_cpu_tensors():
    """There is no CPU path: the resident map must live on the GPU."""
    z = torch.zeros((4, 3))
    with pytest.raises(RuntimeError):
        GaussianMap(z, z, torch.zeros((4, 4)), torch.zeros((4, 1)), z, spatial_order=False)
