"""Drop-in installation into an UNMODIFIED reference checkout.

    import centerpoly_b200.install as cp
    cp.install(reference_src_lib="/path/to/CenterPoly/src/lib")

* DCN: puts the overlay directory ahead of ``src/lib`` on ``sys.path`` so that
  ``models.networks.DCNv2.dcn_v2`` resolves to centerpoly_b200 (see overlay/.../dcn_v2.py).
* decode: rebinds ``polydet_decode`` in ``models.decode`` and in the modules that imported it by name
  (``from models.decode import polydet_decode``: src/lib/detectors/polydet.py:14, src/lib/trains/polydet.py:11).
"""
from __future__ import annotations

import os
import sys

OVERLAY_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "overlay")
REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def install_dcn_overlay(reference_src_lib: str | None = None) -> None:
    for path in (REPO_ROOT, OVERLAY_DIR):
        if path in sys.path:
            sys.path.remove(path)
    # overlay first, then (optionally) the reference, then everything else
    sys.path.insert(0, OVERLAY_DIR)
    sys.path.insert(1, REPO_ROOT)
    if reference_src_lib:
        if reference_src_lib in sys.path:
            sys.path.remove(reference_src_lib)
        sys.path.insert(2, reference_src_lib)


def install_decode() -> None:
    from .decode import polydet_decode
    import importlib

    decode_mod = importlib.import_module("models.decode")
    decode_mod.polydet_decode = polydet_decode
    for name in ("detectors.polydet", "trains.polydet"):
        mod = sys.modules.get(name)
        if mod is not None and hasattr(mod, "polydet_decode"):
            mod.polydet_decode = polydet_decode


def install_poly_loss() -> None:
    """Rebind ``PolyLoss`` where the reference's trainer looks it up (``from models.losses import ... PolyLoss``,
    src/lib/trains/polydet.py:10-12; constructed at :32).  Call it BEFORE ``PolydetLoss`` is instantiated.  Importing
    ``models.losses`` needs the reference's own dependencies (matplotlib, seaborn at losses.py:18-19)."""
    from .poly_loss import PolyLoss
    import importlib

    importlib.import_module("models.losses").PolyLoss = PolyLoss
    mod = sys.modules.get("trains.polydet")
    if mod is not None and hasattr(mod, "PolyLoss"):
        mod.PolyLoss = PolyLoss


def install(reference_src_lib: str | None = None) -> None:
    install_dcn_overlay(reference_src_lib)
    if reference_src_lib:
        install_decode()
