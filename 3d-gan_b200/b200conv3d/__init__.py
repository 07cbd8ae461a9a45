"""b200conv3d — B200-native (sm_100a) conv3d training hot path for 3D-GAN.

Public API: the drop-in layers in :mod:`b200conv3d.nn`, the functional ops in
:mod:`b200conv3d.functional`, ``install()/convert()/fuse()`` in
:mod:`b200conv3d.patch` and the data-parallel gradient reducer in
:mod:`b200conv3d.parallel`.
"""
from . import _lib, functional, nn, patch, parallel, dcgan, train, fid_resnet, biggan, stylegan_ops, stylegan  # noqa: F401
from .patch import convert, fuse, install, prepare, uninstall  # noqa: F401

__version__ = "0.1.0"
