"""Training engines: ``NativeTrainer`` (hand-written sm_100a kernels, the product), ``HybridTrainer``
(same kernels for FPN / heads / losses / optimizer / allreduce, autograd for a non-ResNet backbone), ``SpecTrainer``
(PyTorch autograd + KerasAdam + Horovod-style optimizer; CPU capable), ``LibraryTrainer`` (cuDNN +
DDP/NCCL measured baseline)."""


def create_trainer(model, device, lr=1e-5, clipnorm=0.001, batch=1, image_hw=(800, 1333), freeze_backbone=False,
                   engine="auto"):
    """Pick the engine for ``device``: native on CUDA, spec elsewhere."""
    import torch
    from .. import parallel as hvd
    device = torch.device(device)
    if engine == "auto":
        engine = "native" if device.type == "cuda" else "spec"
    if engine == "native":
        from .trainer import NativeTrainer
        from ..models.resnet import ResNet
        cls = NativeTrainer
        if not isinstance(model.backbone, ResNet):
            from .hybrid import HybridTrainer
            cls = HybridTrainer
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
