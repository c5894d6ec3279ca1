"""The launch chain of lgk_post_physics_step has three forms: two launches with hand-over flags
(default from 12 288 envs), two launches with a grid-wide wait (default below), and the three-launch chain of earlier
in the round (LGK_CHAIN=3, kept for A/B measurements).  LGK_CHAIN is read once per process, so each forced form runs
a slice of the parity suite in a process of its own: golden replays of every scenario, the oracle at the BASELINE
sizes, ragged env counts, the python-reward-term path and graph capture."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SLICE = ("test_cuda_path_matches_reference_golden or test_step_matches_oracle_at_the_baseline_sizes or "
         "test_ragged_env_counts or test_python_reward_terms_of_a_subclass_enter_the_fused_kernel or "
         "test_step_is_graph_capturable_and_sync_free or test_full_size_properties_with_in_kernel_rng")


@pytest.mark.gpu
@pytest.mark.parametrize("chain", ["3", "2w", "2f"])
def test_parity_slice_with_the_chain_forced(chain):
    env = dict(os.environ, LGK_CHAIN=chain)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-m", "gpu", "-q",
                        "-x", "-k", SLICE, "-p", "no:cacheprovider"], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=900)
    tail = (r.stdout or "")[-1500:] + (r.stderr or "")[-500:]
    assert r.returncode == 0, f"LGK_CHAIN={chain}: {tail}"
    assert " passed" in r.stdout and "failed" not in r.stdout, tail
