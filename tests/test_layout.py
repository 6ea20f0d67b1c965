"""CPU: host-side contract of the drop-in surface."""
import os
import re

import pytest
import torch

import common
import s2p_b200


def test_product_never_imports_oracle(repo_root):
    pat = re.compile(r"^\s*(from|import)\s+(oracle|s2p_oracle)\b", re.M)
    for dirpath, _, files in os.walk(os.path.join(repo_root, "s2p_b200")):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(dirpath, f)).read()
                assert not pat.search(txt), f"{f} imports the oracle"


@pytest.mark.parametrize("env", ["cheetah", "walker"])
def test_state_dict_is_interchangeable_with_oracle(env):
    ora = common.build_oracle(env, "xavier")
    net = s2p_b200.S2PGeneratorB200(s2p_b200.make_options(env_type=env))
    a, b = ora.state_dict(), net.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(a[k].shape == b[k].shape and a[k].dtype == b[k].dtype for k in a)
    net.load_state_dict(a)
    ora.load_state_dict(net.state_dict())


def test_no_cpu_fallback():
    opt = s2p_b200.make_options(gpu_ids="-1")
    assert opt.gpu_ids == []
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        s2p_b200.define_G(opt)
    net = s2p_b200.S2PGeneratorB200(s2p_b200.make_options())
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net(torch.zeros(1, 3, 100, 100), torch.zeros(1, 17))
    with pytest.raises(ValueError):
        s2p_b200.define_G(s2p_b200.make_options(netG="spade"))


def test_option_names_follow_reference_cli():
    opt = s2p_b200.make_options(["--env_type=walker", "--netG=s2p", "--batchSize=16", "--gpu_ids=0,1",
                                 "--start_idx=3", "--seq_len=5", "--dataroot=./datasets"])
    assert (opt.env_type, opt.state_dim, opt.batchSize, opt.gpu_ids, opt.start_idx, opt.seq_len) == \
        ("walker", 24, 16, [0, 1], 3, 5)
