"""centerpoly_b200 — B200 (sm_100a) implementation of CenterPoly v2's data-parallel hot path:
the DCNv2 modulated deformable convolution of the DLA-34 neck (forward + backward) and the fused
``polydet_decode``.  Host code is Python/PyTorch for memory and streams only; the arithmetic lives in
``libcpoly.so`` (hand-written CUDA behind the C ABI of ``include/cpoly.h``).
"""
from .dcn_v2 import DCN, DCNv2, _DCNv2, dcn_v2_conv, invalidate_packed_weights, set_dcn_path  # noqa: F401
from .decode import polydet_decode, polydet_decode_from_logits  # noqa: F401
from .raster import CITYSCAPES_CLASSES, RasterResult, rasterise_instances  # noqa: F401
from .post_process import batch_inverse_affine, group_by_class, inverse_affine, polydet_post_process  # noqa: F401

__version__ = "0.1.0"
from .poly_loss import IOU_STATUS, PolyLoss, poly_iou_loss, poly_order_loss  # noqa: F401
from .batchnorm import BatchNorm2dTrain, batch_norm_train, use_training_batchnorm  # noqa: F401
from .raster_fill import FillRasterResult, rasterise_fill_instances, write_idd, write_kitti  # noqa: F401
