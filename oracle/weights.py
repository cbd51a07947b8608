"""TEST INFRASTRUCTURE ONLY -- the synthetic state_dict generators the golden vectors were made with.

They live in matcha_b200/synthetic.py (the benchmark needs random-init weights too and the product may not import
oracle/); this module re-exports them so the oracle-side scripts keep one spelling."""
from matcha_b200.synthetic import (estimator_state_dict, generator_state_dict, prefix_mask,  # noqa: F401
                                   text_encoder_state_dict)
