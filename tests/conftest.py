import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA B200 (sm_100) device; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def cfg():
    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    return Qwen3TTSTokenizerV2DecoderConfig()


@pytest.fixture(scope="session")
def sd(cfg):
    from qwen3_tts_b200.synthetic import make_decoder_state_dict
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    return make_decoder_state_dict(cfg, seed=0)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def gpu_model(sd):
    """One decoder handle (fp16 parity mode) shared by the -m gpu tests."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qwen3_tts_b200 import synthetic_model
    return synthetic_model(state_dict=sd, device="cuda:0", compute_mode="fp16")
