"""Overlay module: with ``<repo>/centerpoly_b200/overlay`` AHEAD of the reference's ``src/lib`` on
``sys.path``, the reference's own ``from .DCNv2.dcn_v2 import DCN``
(src/lib/models/networks/pose_dla_dcn.py:16, resnet_dcn.py:18) resolves here, with the reference tree
untouched: ``models`` and ``models.networks`` are PEP-420 namespace packages in the reference (no
``__init__.py``), and its own ``networks/DCNv2`` directory is an empty, un-vendored submodule.
"""
from centerpoly_b200.dcn_v2 import DCN, DCNv2, _DCNv2, dcn_v2_conv  # noqa: F401
