"""Drop-in replacement for the reference's find_len_scales module (same names, same signatures).
The fork pool of the reference (find_len_scales.py:73-77) is gone: a swarm evaluation is one batched
GPU launch, and nothing forks after CUDA initialisation."""
import _bootstrap  # noqa: F401
import numpy as np

from gaussian_process_b200 import pso

sigma_to_percent = pso.sigma_to_percent
evaluate_loss_batch = pso.evaluate_loss_batch


def len_scale_opt(x_known, y_known, e_known, PSO_progress, **kwargs):
    """PSO length-scale search (reference :22-150).  Optional keywords (num_particles, max_iter,
    patience, max_points, lhs_random_state) default to the reference's hard-coded values."""
    return pso.len_scale_opt(x_known, y_known, e_known, PSO_progress, **kwargs)


def evaluate_loss(lengths, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds):
    """Positive calibration loss of one length-scale vector (reference :181-182)."""
    pos = np.asarray(lengths, dtype=np.float64)[None, :]
    return float(pso.evaluate_loss_batch(pos, x_known, y_known, e_known, sigma_vals, expected_percents,
                                         lower_bounds, upper_bounds)[0])


def wass_loss(ls, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds):
    """Negative of evaluate_loss (reference :154-177); -1e13 outside the bounds."""
    return -evaluate_loss(ls, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds)


def evaluate_loss_helper(args):
    """Tuple-unpacking form kept for callers of the reference's pool helper (:186-188)."""
    p, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds = args
    return evaluate_loss(p, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds)
