"""Training engines: ``NativeTrainer`` (hand-written sm_100a kernels, ResNet: the product), ``VGGTrainer`` /
``MobileNetTrainer`` / ``DenseNetTrainer`` (the other backbone families on the same and on their own layer kernels),
``HybridTrainer`` (their base class; alone: native FPN / heads / losses / optimizer / allreduce + an autograd
backbone, the fallback for unknown backbone modules), ``SpecTrainer``
(PyTorch autograd + KerasAdam + Horovod-style optimizer; CPU capable), ``LibraryTrainer`` (cuDNN +
DDP/NCCL measured baseline)."""


def native_engine_class(model):
    """Every backbone family runs entirely on the native kernels; HybridTrainer (autograd backbone) remains as the
    fallback for backbone classes this function does not know."""
    from ..models.resnet import ResNet
    from ..models.vgg import VGG
    if isinstance(model.backbone, ResNet):
        from .trainer import NativeTrainer
        return NativeTrainer
    if isinstance(model.backbone, VGG):
        from .vgg_native import VGGTrainer
        return VGGTrainer
    from ..models.mobilenet import MobileNet
    if isinstance(model.backbone, MobileNet):
        from .mobilenet_native import MobileNetTrainer
        return MobileNetTrainer
    from ..models.densenet import DenseNet
    if isinstance(model.backbone, DenseNet):
        from .densenet_native import DenseNetTrainer
        return DenseNetTrainer
    from .hybrid import HybridTrainer
    return HybridTrainer


def create_trainer(model, device, lr=1e-5, clipnorm=0.001, batch=1, image_hw=(800, 1333), freeze_backbone=False,
                   engine="auto"):
    """Pick the engine for ``device``: native on CUDA, spec elsewhere."""
    import torch
    from .. import parallel as hvd
    device = torch.device(device)
    if engine == "auto":
        engine = "native" if device.type == "cuda" else "spec"
    if engine == "native":
        cls = native_engine_class(model)
        comm = None
        if hvd.is_initialized() and hvd.size() > 1:
            from ..parallel.p2p import P2PComm
            comm = P2PComm(hvd.rank(), hvd.size(), device, hvd._allgather_bytes)
        return cls(model=model, device=device, lr=lr, clipnorm=clipnorm, batch=batch, image_hw=image_hw,
                             freeze_backbone=freeze_backbone, comm=comm, shape_bucket=64)
    from .spec import SpecTrainer
    if freeze_backbone:
        model.freeze_backbone()
    return SpecTrainer(model, device=device, lr=lr, clipnorm=clipnorm)
