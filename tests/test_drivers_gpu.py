"""GPU: the reference's documented entry points (README.md:33,59) re-created at the repo root: train.py for a few
iterations with checkpoint + resume, then simple_test.py rolling out from the saved generator."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dataset(path, n=24):
    from s2p_b200.synthesis import save_dataset_dir
    rs = np.random.RandomState(0)
    save_dataset_dir(path, {"image_observations": rs.randint(0, 256, size=(n, 100, 100, 3)).astype(np.uint8),
                            "image_observations_tp1": rs.randint(0, 256, size=(n, 100, 100, 3)).astype(np.uint8),
                            "next_observations": rs.randn(n, 17).astype(np.float32)})


def test_train_resume_and_simple_test(tmp_path, repo_root):
    import simple_test
    import train
    ds, ck, res = str(tmp_path / "cheetah"), str(tmp_path / "ck"), str(tmp_path / "res")
    _dataset(ds)
    common = [f"--dataroot={ds}", "--env_type=cheetah", "--netG=s2p", "--gpu_ids=0", "--init_type=scaled", f"--checkpoints_dir={ck}"]
    it = train.main(common + ["--batchSize=4", "--niter=1", "--print_freq=2"])
    assert it == 6                                                      # 24 rows / batch 4
    assert os.path.exists(os.path.join(ck, "cheetah_1.pth")) and os.path.exists(os.path.join(ck, "cheetah_latest_trainer.pth"))
    state = torch.load(os.path.join(ck, "cheetah_latest_trainer.pth"), map_location="cpu")
    assert state["epoch"] == 1 and int(state["trainer"]["optG"]["step"]) == 6
    it2 = train.main(common + ["--batchSize=4", "--niter=2", "--continue_train"])
    assert it2 == 12                                                    # resumed at iteration 6, one more epoch
    state2 = torch.load(os.path.join(ck, "cheetah_latest_trainer.pth"), map_location="cpu")
    assert int(state2["trainer"]["optG"]["step"]) == 12 and os.path.exists(os.path.join(ck, "cheetah_2.pth"))
    out = simple_test.main(common + ["--start_idx=3", "--seq_len=5", "--which_epoch=2", f"--results_dir={res}"])
    assert out.shape == (5, 100, 100, 3) and out.dtype == np.uint8
    saved = np.load(os.path.join(res, "cheetah_3", "frames.npy"))
    assert np.array_equal(saved, out) and os.path.exists(os.path.join(res, "cheetah_3", "frame_05.png"))
    with pytest.raises(RuntimeError):
        simple_test.main(common[:3] + ["--gpu_ids=-1"])
