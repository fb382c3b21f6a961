"""Host-side logic that needs no GPU: validation metrics, early stopping, node-loader cycling."""
import numpy as np


def test_eval_metrics_match_sklearn_sweeps():
    import sklearn.metrics as skm
    from nichecompass_b200.train.metrics import eval_metrics
    rng = np.random.default_rng(0)
    labels = (rng.random(500) < 0.5).astype(np.float64)
    probs = np.clip(0.35 * labels + rng.random(500) * 0.7, 0, 1)
    out = eval_metrics(probs, labels)
    best_acc, best_f1, best_thr = 0, 0, 0
    for t in np.arange(0.01, 1, 0.005):
        pred = (probs > t).astype(int)
        a = skm.accuracy_score(labels, pred)
        if a > best_acc:
            best_acc, best_thr = a, t
        best_f1 = max(best_f1, skm.f1_score(labels, pred))
    assert abs(out["best_acc_score"] - best_acc) < 1e-12 and abs(out["best_acc_threshold"] - best_thr) < 1e-12
    assert abs(out["best_f1_score"] - best_f1) < 1e-12


def test_early_stopping_state_machine():
    from nichecompass_b200.train.utils import EarlyStopping
    es = EarlyStopping(patience=3, lr_patience=2, lr_factor=0.1)
    seq = [1.0, 0.9, 0.95, 0.96, 0.97, 0.98]
    out = [es.step(v) for v in seq]
    # epochs < patience never stop; then 2 non-improving epochs reduce the lr, 3 stop
    assert out[0] == (True, False) and out[1] == (True, False)
    assert out[3] == (True, True)                  # epoch 4: 2 epochs without improvement -> reduce lr
    assert out[4] == (False, False)                # epoch 5: 3 epochs without improvement -> stop
    assert es.update_state(0.5) and not es.update_state(0.6)


def test_cycle_iterable_restarts():
    from nichecompass_b200.train.utils import _cycle_iterable

    class L:
        def __init__(self):
            self.starts = 0

        def __iter__(self):
            self.starts += 1
            return iter([1, 2])
    ld = L()
    it = _cycle_iterable(ld)
    assert [next(it) for _ in range(5)] == [1, 2, 1, 2, 1] and ld.starts == 3


def test_complement_index_lists():
    from nichecompass_b200.models.nichecompass import _complement
    assert _complement([0, 2, 5], 7) == [1, 3, 4, 6]
    assert _complement(np.array([], dtype=int), 3) == [0, 1, 2]
