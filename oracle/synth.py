"""The synthetic workload generator lives in the product (vae_npvc_b200/synth.py: bench.py's own arm feeds the CUDA path
from it); the oracle and the tests use the SAME functions so both sides of every parity check see identical inputs."""
from vae_npvc_b200.synth import *  # noqa: F401,F403
from vae_npvc_b200.synth import (arch_config, cdvae_batch, cdvae_config, cdvae_state_dict, conv_macs_per_frame,  # noqa: F401
                                 small_config, synthetic_batch, synthetic_state_dict)
