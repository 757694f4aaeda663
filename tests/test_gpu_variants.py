"""GPU parity of the code paths that are selected per process by an environment switch (read once by the library):
the fused MLP kernel of the coarse encoder layer (OPB200_MLP_FUSED=1 forces it -- by default it runs from four tiles per
CTA on, i.e. not at most test shapes -- and =0 forces the two-launch path it replaced), the fused ATTN + MERGE + LayerNorm1 kernel of the coarse level
(OPB200_ATTN_MERGE_FUSED, same convention) and the CUDA-core KV kernel that
the tensor-core one replaced (OPB200_KV_TC=0); the per-tap convolution kernels that the halo kernels replaced for 3x3 stride 1
(OPB200_CONV_HALO=0).  Each variant runs the encoder / model parity tests in a child process."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("env", [{"OPB200_MLP_FUSED": "1"}, {"OPB200_MLP_FUSED": "0"}, {"OPB200_KV_TC": "0"},
                                 {"OPB200_ATTN_MERGE_FUSED": "1", "OPB200_MLP_FUSED": "1"}, {"OPB200_ATTN_MERGE_FUSED": "0"},
                                 {"OPB200_MLP_FUSED": "1", "OPB200_MLP_PHASE0": "0"}],
                         ids=["mlp_fused", "mlp_two_launches", "kv_cuda_cores", "attn_merge_fused", "attn_merge_separate",
                              "mlp_fused_without_phase0"])
def test_encoder_variant_passes_the_parity_tests(env):
    res = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                          os.path.join(REPO, "tests", "test_gpu_transformer.py"),
                          os.path.join(REPO, "tests", "test_gpu_models.py")],
                         cwd=REPO, env=dict(os.environ, **env), capture_output=True, text=True, timeout=900)
    tail = (res.stdout + res.stderr)[-1500:]
    assert res.returncode == 0, tail
    assert " passed" in res.stdout and " failed" not in res.stdout, tail


def test_per_tap_convolution_kernels_pass_the_backbone_tests():
    res = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                          os.path.join(REPO, "tests", "test_gpu_conv.py"), os.path.join(REPO, "tests", "test_gpu_fine.py")],
                         cwd=REPO, env=dict(os.environ, OPB200_CONV_HALO="0"), capture_output=True, text=True, timeout=900)
    tail = (res.stdout + res.stderr)[-1500:]
    assert res.returncode == 0, tail
    assert " passed" in res.stdout and " failed" not in res.stdout, tail


def test_sweep_pair_mode_passes_the_coarse_tests():
    """OPB200_SIM_PAIR=1: pairs of tiles on 128 x 256 MMAs in the CoarseMatching sweep (measured no faster, kept selectable)."""
    res = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                          os.path.join(REPO, "tests", "test_gpu_coarse.py")],
                         cwd=REPO, env=dict(os.environ, OPB200_SIM_PAIR="1"), capture_output=True, text=True, timeout=900)
    tail = (res.stdout + res.stderr)[-1500:]
    assert res.returncode == 0, tail
    assert " passed" in res.stdout and " failed" not in res.stdout, tail
