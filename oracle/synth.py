"""The synthetic workload generator lives in the product (vae_npvc_b200/synth.py: bench.py feeds the CUDA path from
it); the oracle side and the tests import the SAME functions through this module so both sides of every parity check
see identical configs, weights and batches.  TEST INFRASTRUCTURE (see oracle/__init__.py)."""
from vae_npvc_b200.synth import (arch_config, small_config, synthetic_batch,  # noqa: F401
                                 synthetic_state_dict)
