"""Drop-in replacement for the reference's calc_prob_surf module (same names, same signatures,
same CSV), backed by the fused B200 surface kernel."""
import _bootstrap  # noqa: F401

from gaussian_process_b200 import surface

sum_gaussians = surface.sum_gaussians
generate_prob_surf = surface.generate_prob_surf
