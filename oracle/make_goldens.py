"""Generate ``tests/golden/*.npz`` by running the REAL reference from ``/root/reference``.

Run in the build container only (``python -m oracle.make_goldens``); the GPU box has no
``/root/reference`` and only ever reads the committed ``.npz`` files.

How the reference is loaded:

* ``polyaxon_schemas`` is not installed, so ``schemas.hptuning`` is served by the stand-in in
  ``polyaxon_b200/schemas/hptuning.py`` (config plumbing only).
* ``space.py``, ``optimizer.py``, ``utils.py`` are executed unmodified from
  ``/root/reference/polyaxon/hpsearch/search_managers/``.  The package ``__init__`` files are not
  executed because they pull in marshmallow / the sibling search algorithms.
* ``acquisition_function.py`` gets the 3-line scipy>=1.5 compatibility patch, applied to the source
  text in memory: ``x_try.reshape(1, -1)`` as ``x0`` becomes ``x_try.ravel()`` and the two
  ``res.fun[0]`` become ``np.ravel(res.fun)[0]``.  Nothing is written under ``/root/reference``.

Each golden holds the inputs and every staged intermediate of one ``BOOptimizer.get_suggestion()``
call (theta*, LML, alpha, diag(L), posterior mean/std and acquisition on a candidate block, argmax,
per-seed refinement results, final x*), plus LML/gradient at fixed thetas from sklearn itself.
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np

REF = '/root/reference/polyaxon'
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), 'tests', 'golden')


def load_reference():
    sys.path.insert(0, os.path.dirname(HERE))
    from polyaxon_b200.schemas import hptuning as standin

    schemas_pkg = types.ModuleType('schemas')
    schemas_pkg.__path__ = []
    sys.modules['schemas'] = schemas_pkg
    sys.modules['schemas.hptuning'] = standin
    for name, rel in (('hpsearch', 'hpsearch'),
                      ('hpsearch.search_managers', 'hpsearch/search_managers'),
                      ('hpsearch.search_managers.bayesian_optimization',
                       'hpsearch/search_managers/bayesian_optimization')):
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(REF, rel)]
        sys.modules[name] = pkg

    def load(name, rel, patch=None):
        path = os.path.join(REF, rel)
        src = open(path).read()
        if patch:
            for old, new, count in patch:
                assert src.count(old) == count, (old, src.count(old))
                src = src.replace(old, new)
        mod = types.ModuleType(name)
        mod.__file__ = path
        sys.modules[name] = mod
        exec(compile(src, path, 'exec'), mod.__dict__)
        return mod

    load('hpsearch.search_managers.utils', 'hpsearch/search_managers/utils.py')
    load('hpsearch.search_managers.base', 'hpsearch/search_managers/base.py')
    load('hpsearch.search_managers.bayesian_optimization.space',
         'hpsearch/search_managers/bayesian_optimization/space.py')
    load('hpsearch.search_managers.bayesian_optimization.acquisition_function',
         'hpsearch/search_managers/bayesian_optimization/acquisition_function.py',
         patch=[('                           x_try.reshape(1, -1),\n',
                 '                           x_try.ravel(),\n', 1),
                ('res.fun[0]', 'np.ravel(res.fun)[0]', 2)])
    opt = load('hpsearch.search_managers.bayesian_optimization.optimizer',
               'hpsearch/search_managers/bayesian_optimization/optimizer.py')
    mgr = load('hpsearch.search_managers.bayesian_optimization.manager',
               'hpsearch/search_managers/bayesian_optimization/manager.py')
    return standin, opt, mgr


def hp_config(standin, d, kernel, nu, acq, kappa, eps, n_restarts, n_warmup, n_iter, seed,
              length_scale=1.0, matrix=None, optimization='maximize'):
    uf = {'acquisition_function': acq,
          'gaussian_process': {'kernel': kernel, 'length_scale': length_scale, 'nu': nu,
                               'n_restarts_optimizer': n_restarts},
          'n_warmup': n_warmup, 'n_iter': n_iter}
    if acq == 'ucb':
        uf['kappa'] = kappa
    else:
        uf['eps'] = eps
    if matrix is None:
        matrix = {'x%03d' % i: {'uniform': [0, 1]} for i in range(d)}
    return standin.HPTuningConfig.from_dict({
        'seed': seed, 'concurrency': 2,
        'bo': {'n_iterations': 5, 'n_initial_trials': 5,
               'metric': {'name': 'm', 'optimization': optimization},
               'utility_function': uf},
        'matrix': matrix})


def run_case(name, standin, opt_mod, cfg, configs, metrics, fixed_thetas, n_block):
    """One BOOptimizer.get_suggestion() of the real reference, instrumented from outside."""
    from sklearn.gaussian_process import GaussianProcessRegressor
    optimizer = opt_mod.BOOptimizer(hptuning_config=cfg)
    optimizer.add_observations(configs=configs, metrics=metrics)
    uf = optimizer.utility_function
    rec = {}

    # record what the refinement loop sees, without touching the reference's code
    import scipy.optimize as so
    acq_mod = sys.modules['hpsearch.search_managers.bayesian_optimization.acquisition_function']
    real_minimize = so.minimize
    runs = []

    def spy_minimize(fun, x0, **kw):
        res = real_minimize(fun, x0, **kw)
        runs.append((np.array(x0, dtype=float).ravel(), res.x.copy(),
                     -float(np.ravel(res.fun)[0]), bool(res.success), int(res.nfev)))
        return res
    acq_mod.minimize = spy_minimize

    rng_state = uf.random_generator.get_state()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        x_star = optimizer._maximize()
        suggestion = optimizer.space.get_suggestion(x_star)
    acq_mod.minimize = real_minimize

    gp = uf.gaussian_process
    X = np.asarray(optimizer.space.x, dtype=np.float64)
    y = np.asarray(optimizer.space.y, dtype=np.float64)
    bounds = np.asarray(optimizer.space.bounds, dtype=np.float64)
    d = bounds.shape[0]
    R = 1 + cfg.bo.utility_function.gaussian_process.n_restarts_optimizer

    # replay the RNG stream in the contract order: restarts -> x_tries -> x_seeds
    rs = np.random.RandomState()
    rs.set_state(rng_state)
    lo, hi = np.log(1e-5), np.log(1e5)
    theta_inits = np.array([rs.uniform(lo, hi) for _ in range(R - 1)])
    x_tries = rs.uniform(bounds[:, 0], bounds[:, 1], size=(optimizer.n_warmup, d))
    x_seeds = rs.uniform(bounds[:, 0], bounds[:, 1], size=(optimizer.n_iter, d))
    y_max = y.max()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        ys = uf.compute(x_tries, y_max=y_max)
        blk = x_tries[:n_block]
        mean, std = gp.predict(blk, return_std=True)
    assert np.allclose(np.array([r[0] for r in runs]), x_seeds), 'RNG contract broken'

    # per-restart optima straight from sklearn's own optimiser
    per_restart = []
    for t0 in np.concatenate([[np.log(cfg.bo.utility_function.gaussian_process.length_scale)],
                              theta_inits]):
        g2 = GaussianProcessRegressor(kernel=gp.kernel, optimizer=gp.optimizer)
        g2.kernel_ = gp.kernel_.clone_with_theta(np.array([t0]))
        g2.X_train_, g2.y_train_ = X, y
        g2.alpha = gp.alpha
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            th, fmin = g2._constrained_optimization(
                lambda th, eval_gradient=True: tuple(
                    -v for v in g2.log_marginal_likelihood(th, eval_gradient=True, clone_kernel=False)),
                np.array([t0]), gp.kernel_.bounds)
        per_restart.append((t0, float(th[0]), float(-fmin)))

    lml_fixed, grad_fixed = [], []
    for th in fixed_thetas:
        l, g = gp.log_marginal_likelihood(np.array([th]), eval_gradient=True)
        lml_fixed.append(l)
        grad_fixed.append(g[0])

    kern = cfg.bo.utility_function.gaussian_process.kernel
    rec.update(
        name=name, X=X, y=y, bounds=bounds, seed=cfg.seed,
        kernel=0 if kern == 'rbf' else 1,
        nu=float(cfg.bo.utility_function.gaussian_process.nu),
        length_scale=float(cfg.bo.utility_function.gaussian_process.length_scale),
        acq={'ucb': 0, 'ei': 1, 'poi': 2}[uf.acquisition_function],
        kappa=np.nan if uf.kappa is None else float(uf.kappa),
        eps=np.nan if uf.eps is None else float(uf.eps),
        n_restarts=R - 1, n_warmup=optimizer.n_warmup, n_iter=optimizer.n_iter,
        theta_inits=theta_inits, per_restart=np.array(per_restart),
        theta=float(gp.kernel_.theta[0]), lml=float(gp.log_marginal_likelihood_value_),
        alpha=gp.alpha_, L_diag=np.diag(gp.L_).copy(),
        L=gp.L_ if X.shape[0] <= 64 else np.zeros((0, 0)),
        cond=float(np.linalg.cond(gp.L_ @ gp.L_.T)),
        fixed_thetas=np.array(fixed_thetas), lml_fixed=np.array(lml_fixed),
        grad_fixed=np.array(grad_fixed),
        x_tries_head=x_tries[:n_block], x_tries_sum=float(x_tries.sum()),
        mean=mean, std=std, ys_head=ys[:n_block], y_max=float(y_max),
        argmax=int(ys.argmax()), warm_acq=float(ys.max()), x_warm=x_tries[ys.argmax()],
        x_seeds=x_seeds,
        ref_x=np.array([r[1] for r in runs]), ref_acq=np.array([r[2] for r in runs]),
        ref_success=np.array([r[3] for r in runs]), ref_nfev=np.array([r[4] for r in runs]),
        x_star=np.asarray(x_star, dtype=np.float64),
        suggestion_keys=np.array(sorted(suggestion.keys())),
        suggestion_repr=np.array([repr(suggestion[k]) for k in sorted(suggestion.keys())]),
    )
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **rec)
    print('%-28s n=%-5d d=%-3d theta*=%.6f lml=%.6f cond=%.2e argmax=%d warm=%.6g x*[0]=%.6g' % (
        name, X.shape[0], d, rec['theta'], rec['lml'], rec['cond'], rec['argmax'],
        rec['warm_acq'], rec['x_star'][0]))


def synthetic_configs(n, d, seed=1234):
    from oracle.reference_path import synthetic_problem
    X, y, _ = synthetic_problem(n, d, seed)
    keys = ['x%03d' % i for i in range(d)]
    return [dict(zip(keys, row.tolist())) for row in X], y.tolist()


# the 14-point observation set of test_search_managers.py:894-912 (features 1..3, minimise)
REF14_X = [(1, 1, 1), (1, 1.2, 1), (1, 1, 1), (1, 1.11, 1), (1, 1.1, 1), (1, 1.21, 1), (2, 2, 2),
           (3, 2, 2), (2, 1.8, 3), (3, 2, 3), (2, 2, 2), (3, 2, 2), (2, 1.8, 3), (3, 2, 3)]
REF14_Y = [0, 1.1, 0.1, 0.1, 1.09, 0.4, 100, 200, 200, 300, 110, 210, 210, 310]
# the 10-point set of test_search_managers.py:970-1063 (learning_rate, activation, dropout)
REF10 = [(0.004544653508229265, 'sigmoid', 0.3), (0.005615296199690899, 'sigmoid', 0.3),
         (0.008784330869587902, 'sigmoid', 0.25), (0.0058591075447430065, 'sigmoid', 0.3),
         (0.007464080062927171, 'sigmoid', 0.25), (0.0024763129571936738, 'relu', 0.3),
         (0.0074881581817925705, 'sigmoid', 0.3), (0.003360405779075163, 'relu', 0.3),
         (0.009916904455792564, 'sigmoid', 0.25), (0.000881723263162717, 'sigmoid', 0.3)]
REF10_Y = [2.3018131256103516, 2.302884340286255, 2.3071441650390625, 2.3034636974334717,
           2.301487922668457, 0.05087224021553993, 2.3032383918762207, 0.06383182853460312,
           2.3120572566986084, 0.7617478370666504]


def main(only=None):
    standin, opt_mod, _ = load_reference()
    thetas = [-2.0, -1.0, 0.0, 0.7, 1.5, 3.0]
    if only:
        real_run_case = run_case

        def filtered(name, *args, **kw):
            if name in only:
                real_run_case(name, *args, **kw)
        globals()['run_case'] = filtered

    # C1: d=2, Matern 2.5, UCB kappa=2.576, n=15 (5 initial + 10 iterations)
    c, m = synthetic_configs(15, 2)
    run_case('c1_d2_n15_m25_ucb', standin, opt_mod,
             hp_config(standin, 2, 'matern', 2.5, 'ucb', 2.576, None, 0, 20000, 10, 7),
             c, m, thetas, 2048)
    # C2: d=8 mixed (3 continuous + 2 discrete + one 3-way categorical), RBF, EI eps=0, n=256
    rs = np.random.RandomState(99)
    mixed = {'a_lr': {'uniform': [0, 1]}, 'b_mom': {'uniform': [0, 1]}, 'c_wd': {'uniform': [0, 1]},
             'd_layers': {'range': [1, 9, 1]}, 'e_width': {'linspace': [0, 1, 11]},
             'f_act': {'values': ['relu', 'tanh', 'gelu']}}
    cfgs, mets = [], []
    for _ in range(256):
        p = {'a_lr': rs.uniform(), 'b_mom': rs.uniform(), 'c_wd': rs.uniform(),
             'd_layers': int(rs.randint(1, 9)), 'e_width': float(rs.choice(np.linspace(0, 1, 11))),
             'f_act': str(rs.choice(['relu', 'tanh', 'gelu']))}
        cfgs.append(p)
        mets.append(float(np.sin(3 * p['a_lr']) + 0.5 * p['b_mom'] - 0.3 * p['c_wd']
                          + 0.05 * p['d_layers'] + 0.2 * p['e_width']
                          + {'relu': 0.0, 'tanh': 0.1, 'gelu': 0.2}[p['f_act']]
                          + 0.01 * rs.randn()))
    run_case('c2_d8_n256_rbf_ei', standin, opt_mod,
             hp_config(standin, 8, 'rbf', 1.5, 'ei', None, 0.0, 0, 8192, 4, 11, matrix=mixed),
             cfgs, mets, thetas, 2048)
    # C3 scaled down: d=32, Matern 1.5, POI eps=0, n=512, 2 restarts
    c, m = synthetic_configs(512, 32)
    run_case('c3_d32_n512_m15_poi', standin, opt_mod,
             hp_config(standin, 32, 'matern', 1.5, 'poi', None, 0.0, 2, 4096, 2, 5),
             c, m, thetas, 1024)
    # C3 at its real shape: d=32, n=2048, Matern 1.5, POI eps=0, 17 optimiser runs, 2^18 candidates swept by the
    # reference's own single compute() call (K* is 2^18 x 2048 doubles = 4.3 GB; ~30 GB peak with sklearn's
    # temporaries, ~10 minutes on 8 cores).  acquisition_function.py:88-92 at the headline size.
    c, m = synthetic_configs(2048, 32)
    run_case('c3_d32_n2048_m15_poi_full', standin, opt_mod,
             hp_config(standin, 32, 'matern', 1.5, 'poi', None, 0.0, 16, 1 << 18, 3, 5),
             c, m, thetas, 4096)
    # C5 member: d=16, n=512 would repeat C3's shape; use n=128, Matern 0.5 + RBF variants instead
    c, m = synthetic_configs(128, 16, seed=1240)
    run_case('c5_d16_n128_m05_ucb', standin, opt_mod,
             hp_config(standin, 16, 'matern', 0.5, 'ucb', 1.5, None, 1, 4096, 3, 21),
             c, m, thetas, 1024)
    run_case('c5_d16_n128_m25_ei', standin, opt_mod,
             hp_config(standin, 16, 'matern', 2.5, 'ei', None, 0.01, 3, 4096, 3, 22),
             c, m, thetas, 1024)
    # the reference's own observation sets (closed-form nu so the device path covers them)
    m3 = {'feature1': {'values': [1, 2, 3]}, 'feature2': {'linspace': [1, 2, 5]},
          'feature3': {'range': [1, 5, 1]}}
    c14 = [dict(zip(('feature1', 'feature2', 'feature3'), r)) for r in REF14_X]
    run_case('ref14_m25_ucb', standin, opt_mod,
             hp_config(standin, 3, 'matern', 2.5, 'ucb', 1.2, None, 0, 4096, 10, 3, matrix=m3,
                       optimization='minimize'),
             c14, REF14_Y, thetas, 1024)
    m10 = {'learning_rate': {'uniform': [0.001, 0.01]}, 'dropout': {'values': [0.25, 0.3]},
           'activation': {'values': ['relu', 'sigmoid']}}
    c10 = [{'learning_rate': a, 'activation': b, 'dropout': c_} for a, b, c_ in REF10]
    run_case('ref10_m15_ucb', standin, opt_mod,
             hp_config(standin, 4, 'matern', 1.5, 'ucb', 2.576, None, 0, 4096, 10, 5, matrix=m10,
                       optimization='minimize'),
             c10, REF10_Y, thetas, 1024)
    # general nu = 1.9 (the value in every reference fixture) on well-conditioned synthetic data
    c, m = synthetic_configs(15, 2)
    run_case('c1_d2_n15_m19_ucb', standin, opt_mod,
             hp_config(standin, 2, 'matern', 1.9, 'ucb', 2.576, None, 0, 8192, 4, 7),
             c, m, thetas, 1024)
    c, m = synthetic_configs(128, 16, seed=1240)
    run_case('c5_d16_n128_m19_ei', standin, opt_mod,
             hp_config(standin, 16, 'matern', 1.9, 'ei', None, 0.01, 1, 2048, 2, 23),
             c, m, thetas, 512)
    # the reference's own two observation sets with the fixture's nu=1.9
    run_case('ref14_m19_ucb', standin, opt_mod,
             hp_config(standin, 3, 'matern', 1.9, 'ucb', 1.2, None, 0, 2048, 5, 3, matrix=m3,
                       optimization='minimize'),
             c14, REF14_Y, thetas, 512)
    run_case('ref10_m19_ucb', standin, opt_mod,
             hp_config(standin, 4, 'matern', 1.9, 'ucb', 2.576, None, 0, 2048, 5, 5, matrix=m10,
                       optimization='minimize'),
             c10, REF10_Y, thetas, 512)


if __name__ == '__main__':
    main(only=set(sys.argv[1:]) or None)
