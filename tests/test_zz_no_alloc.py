"""GPU: s2p_state_encoder_fwd_static -- the state path for callers that cannot allocate (its workspace is a file-scope
device array inside the library) -- called through the C ABI and compared with s2p_state_encoder_fwd (caller-provided
workspace) on the same inputs.  Same kernel, same split-K order: the tests of the existing entry point assert
bit-identity against the pos_encode + linear chain (tests/test_kernels_gpu.py), and the same bar is used here.
Collected last on purpose; passes alone in a fresh process and after the others."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from s2p_b200 import _lib
    return _lib


def _weights(s_dim, width, heads_n, seed):
    g = torch.Generator().manual_seed(seed)
    k0 = s_dim * 21
    layers = []
    for i in range(8):
        k = k0 if i == 0 else width
        layers.append(((torch.randn(width, k, generator=g) / k ** 0.5).cuda(), (torch.randn(width, generator=g) * 0.1).cuda()))
    hw = (torch.randn(heads_n, width, generator=g) / width ** 0.5).cuda()
    hb = (torch.randn(heads_n, generator=g) * 0.1).cuda()
    return layers, hw, hb


def _call(lib, static, state, layers, hw, hb, out):
    L = lib.load()
    b, s = state.shape
    width, n = layers[0][0].shape[0], hw.shape[0]
    k = len(layers)
    wp = (ctypes.c_void_p * k)(*[w.data_ptr() for w, _ in layers])
    bp = (ctypes.c_void_p * k)(*[bias.data_ptr() for _, bias in layers])
    st = torch.cuda.current_stream().cuda_stream
    if static:
        return L.s2p_state_encoder_fwd_static(state.data_ptr(), b, s, 10, wp, bp, k, width, hw.data_ptr(), hb.data_ptr(), n,
                                              out.data_ptr(), out.stride(0), st)
    need = int(L.s2p_state_encoder_workspace_bytes(b, width))
    ws = torch.zeros(need, dtype=torch.uint8, device="cuda")
    rc = L.s2p_state_encoder_fwd(state.data_ptr(), b, s, 10, wp, bp, k, width, hw.data_ptr(), hb.data_ptr(), n,
                                 out.data_ptr(), out.stride(0), ws.data_ptr(), need, st)
    torch.cuda.synchronize()      # ws stays alive until the launch is done
    return rc


@pytest.mark.parametrize("b,s_dim", [(16, 17), (64, 24), (1, 17)])
def test_static_scratch_matches_the_workspace_entry_point_twice(lib, b, s_dim):
    width, heads_n = 512, 8576
    layers, hw, hb = _weights(s_dim, width, heads_n, 7)
    g = torch.Generator().manual_seed(100 + b)
    for call in range(2):          # two calls in a row: the second finds the first one's contents in the shared array
        state = (torch.randn(b, s_dim, generator=g) * 3).cuda()
        want = torch.full((b, heads_n), float("nan"), device="cuda")
        got = torch.full((b, heads_n), float("nan"), device="cuda")
        lib.check(_call(lib, False, state, layers, hw, hb, want), "s2p_state_encoder_fwd")
        n0 = lib.launch_count()
        lib.check(_call(lib, True, state, layers, hw, hb, got), "s2p_state_encoder_fwd_static")
        assert lib.launch_count() - n0 == 1
        torch.cuda.synchronize()
        assert torch.isfinite(want).all()
        assert torch.equal(got, want), (call, float((got - want).abs().max()))


def test_static_scratch_after_a_smaller_and_a_larger_batch(lib):
    # 64 rows fill the array, then 3 rows run over its stale contents; results still equal the workspace path
    layers, hw, hb = _weights(17, 512, 8576, 9)
    g = torch.Generator().manual_seed(5)
    for b in (64, 3, 64):
        state = (torch.randn(b, 17, generator=g) * 3).cuda()
        want = torch.empty(b, 8576, device="cuda")
        got = torch.empty(b, 8576, device="cuda")
        lib.check(_call(lib, False, state, layers, hw, hb, want), "s2p_state_encoder_fwd")
        lib.check(_call(lib, True, state, layers, hw, hb, got), "s2p_state_encoder_fwd_static")
        torch.cuda.synchronize()
        assert torch.equal(got, want), b


def test_static_scratch_rejects_sizes_beyond_its_maximum(lib):
    L = lib.load()
    assert int(L.s2p_state_encoder_static_max_floats()) == 32768
    layers, hw, hb = _weights(17, 512, 64, 3)
    state = torch.zeros(65, 17, device="cuda")           # 65 * 512 floats > 32768
    out = torch.full((65, 64), 7.0, device="cuda")
    n0 = lib.launch_count()
    rc = _call(lib, True, state, layers, hw, hb, out)
    assert rc != 0 and lib.launch_count() == n0
    msg = L.s2p_last_error_string().decode()
    assert "exceeds the built-in scratch" in msg and "32768" in msg
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())                      # nothing was launched
