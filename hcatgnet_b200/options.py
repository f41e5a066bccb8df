"""Default hyper-parameters of the hot path, as an ``argparse.Namespace``.

Only the fields the model constructor reads (reference model/networks.py:13-19,
model/gcn.py:15,48-50) with the defaults of reference options/base_options.py:171-281,
which are also the values stored in every shipped ``model.pth``.
The reference's full CLI (dataset paths, TML options, plotting flags) is out of scope.
"""
import argparse

_DEFAULTS = dict(
    n_classes=1, n_convolutions=2, readout_layers=2, embedding_dim=64, improved=True,
    problem_type="regression", optimizer="Adam", lr=0.01, scheduler="ReduceLROnPlateau",
    step_size=7, gamma=0.7, min_lr=1e-08, batch_size=40, epochs=250, early_stopping=6,
    global_seed=20232023, folds=10,
)


def default_options(**overrides) -> argparse.Namespace:
    unknown = set(overrides) - set(_DEFAULTS)
    if unknown:
        raise TypeError(f"unknown option(s): {sorted(unknown)}")
    return argparse.Namespace(**{**_DEFAULTS, **overrides})
