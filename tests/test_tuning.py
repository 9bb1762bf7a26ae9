"""CMA-ES strategy of the batched controller tuner on analytic objectives; the roll-out half is a gpu test
(tests/test_gpu_env.py::test_batched_tuner_improves_p2p_fitness)."""
import numpy as np

from mujoco_robot_environments_b200.tuning import CMAES


def test_cmaes_minimises_a_shifted_ellipsoid():
    es = CMAES(6, 20, sigma_init=500.0, mean_init=500.0, seed=0)       # the reference's popsize / dims / sigma / mean (:146-153)
    opt = np.array([350.0, 20.0, 500.0, 100.0, 200.0, 30.0])
    scale = np.array([1.0, 10.0, 1.0, 5.0, 2.0, 10.0])
    f = lambda x: (((np.abs(x) - opt) * scale) ** 2).sum(1)
    f0 = None
    for g in range(150):
        x = es.ask()
        assert x.shape == (20, 6)
        fit = f(x)
        f0 = fit.min() if f0 is None else f0
        es.tell(x, fit)
    assert es.best_fitness < 1e-6 * f0
    assert np.abs(np.abs(es.best_member) - opt).max() < 0.1
    assert np.allclose(es.C, es.C.T) and np.linalg.eigvalsh(es.C).min() > 0


def test_cmaes_is_deterministic_per_seed():
    a, b = CMAES(6, 8, 1.0, 0.0, seed=3), CMAES(6, 8, 1.0, 0.0, seed=3)
    for _ in range(5):
        xa, xb = a.ask(), b.ask()
        assert np.array_equal(xa, xb)
        a.tell(xa, (xa ** 2).sum(1)); b.tell(xb, (xb ** 2).sum(1))
    assert a.sigma == b.sigma
