#!/usr/bin/env python
"""Training script -- same CLI as the reference ``train.py`` (flags verbatim, ``train.py:329-380``;
SURVEY.md Appendix A), B200-native underneath.

    python -m batchai_retinanet_horovod_coco_b200.parallel.launcher -np 8 \\
        python train.py --steps 500 --tensorboard-dir ./logs --no-evaluation coco /data/coco

also runs under ``torchrun`` / ``mpirun`` (rank env of either is recognised) and single-process.
Extra flags (not in the reference) are grouped at the end of ``parse_args``.
"""

from __future__ import annotations

import argparse
import os
import sys
import warnings

# Initialise the distributed runtime before anything touches a GPU, like the reference does with
# `import horovod.keras as hvd; hvd.init()` at the very top (train.py:20-21).
if __package__ in (None, ""):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
    __package__ = "batchai_retinanet_horovod_coco_b200.bin"

from .. import parallel as hvd  # noqa: E402

import torch  # noqa: E402

from .. import models  # noqa: E402
from ..callbacks import (BroadcastGlobalVariablesCallback, CallbackList, CocoEval, Evaluate, MetricAverageCallback,  # noqa: E402
                         ModelCheckpoint, ProgbarLogger, RedirectModel, ReduceLROnPlateau, TensorBoard)
from ..engine import create_trainer  # noqa: E402
from ..utils.transform import random_transform_generator  # noqa: E402


def makedirs(path):
    os.makedirs(path, exist_ok=True)


MINIMUM_TORCH = (2, 4)
OPS_ABI = 2


def check_versions(device=None):
    """The reference asserts a minimum Keras version before building anything (``check_keras_version()``,
    train.py:52,393).  The equivalents here: a minimum PyTorch, and -- on the GPU path -- that the in-tree native
    library is the one this Python layer was written against and that the device is a Blackwell part."""
    have = tuple(int(v) for v in torch.__version__.split("+")[0].split(".")[:2])
    assert have >= MINIMUM_TORCH, "You are using PyTorch version {}. The minimum required version is {}.".format(
        torch.__version__, ".".join(map(str, MINIMUM_TORCH)))
    if device is not None and torch.device(device).type == "cuda":
        from ..ops import native
        abi = int(native.require().b200_ops_abi_version())
        assert abi == OPS_ABI, f"libb200ops ABI {abi} does not match this package ({OPS_ABI}); run python build_native.py"
        major, _ = torch.cuda.get_device_capability(device)
        assert major >= 10, "the native engine needs an sm_100a (B200) device; use --cpu for the spec engine elsewhere"


def latest_snapshot(snapshot_path: str):
    """Newest ``checkpoint-NN.h5`` (highest NN) in ``snapshot_path``, or None."""
    import re
    best = None
    if os.path.isdir(snapshot_path):
        for f in os.listdir(snapshot_path):
            m = re.fullmatch(r"checkpoint-(\d+)\.h5", f)
            if m and (best is None or int(m.group(1)) > best[0]):
                best = (int(m.group(1)), os.path.join(snapshot_path, f))
    return best[1] if best else None


def get_device(args):
    """One process per GPU, pinned by local rank (reference ``get_session``, train.py:68-72).  ``--gpu``
    wins over the local rank when given."""
    if args.cpu or not torch.cuda.is_available():
        return torch.device("cpu")
    idx = int(args.gpu) if args.gpu else hvd.local_rank() % max(1, torch.cuda.device_count())
    torch.cuda.set_device(idx)
    return torch.device("cuda", idx)


def model_with_weights(model, weights, skip_mismatch):
    if weights is not None:
        from ..utils.keras_h5 import load_weights
        load_weights(model, weights, by_name=True, skip_mismatch=skip_mismatch)
    return model


def create_models(backbone_retinanet, num_classes, weights, multi_gpu=0, freeze_backbone=False):
    """(model, training_model, prediction_model) -- three views of ONE network (reference train.py:81-107).
    The optimizer/loss 'compile' step lives in the trainer (engine.create_trainer)."""
    modifier = (lambda m: m.freeze_backbone()) if freeze_backbone else None
    model = model_with_weights(backbone_retinanet(num_classes, modifier=modifier), weights=weights, skip_mismatch=True)
    training_model = model
    prediction_model = models.retinanet_bbox(model=model)
    return model, training_model, prediction_model


def create_callbacks(model, training_model, prediction_model, validation_generator, args):
    """Same list, same order, same rank gating as the reference (train.py:110-174)."""
    callbacks = [BroadcastGlobalVariablesCallback(0)]
    if hvd.rank() == 0 and args.snapshots:  # only one worker saves the checkpoint file
        callbacks.append(ModelCheckpoint(os.path.join(args.snapshot_path, "checkpoint-{epoch:02d}.h5")))
    tensorboard_callback = None
    if args.tensorboard_dir and hvd.rank() == 0:
        tensorboard_callback = TensorBoard(log_dir=args.tensorboard_dir, histogram_freq=0, batch_size=args.batch_size,
                                           write_graph=True, write_grads=False, write_images=False)
        callbacks.append(tensorboard_callback)
    if args.evaluation and validation_generator:
        if args.dataset_type == "coco":
            evaluation = CocoEval(validation_generator, tensorboard=tensorboard_callback)
        else:
            evaluation = Evaluate(validation_generator, tensorboard=tensorboard_callback)
        callbacks.append(RedirectModel(evaluation, None))  # None -> the trainer's prediction graph
    callbacks.append(MetricAverageCallback())
    callbacks.append(ReduceLROnPlateau(monitor="loss", factor=0.1, patience=2, verbose=1 if hvd.rank() == 0 else 0,
                                       mode="auto", epsilon=0.0001, cooldown=0, min_lr=0))
    # TensorBoard must see the averaged metrics and lr: move it behind them (Keras passes one shared dict)
    if tensorboard_callback is not None:
        callbacks.remove(tensorboard_callback)
        callbacks.append(tensorboard_callback)
    return callbacks


def create_generators(args):
    """(train_generator, validation_generator) for one of five dataset types (train.py:177-297)."""
    if args.random_transform:
        transform_generator = random_transform_generator(
            min_rotation=-0.1, max_rotation=0.1, min_translation=(-0.1, -0.1), max_translation=(0.1, 0.1),
            min_shear=-0.1, max_shear=0.1, min_scaling=(0.9, 0.9), max_scaling=(1.1, 1.1), flip_x_chance=0.5,
            flip_y_chance=0.5)
    else:
        transform_generator = random_transform_generator(flip_x_chance=0.5)
    common = dict(batch_size=args.batch_size, image_min_side=args.image_min_side, image_max_side=args.image_max_side)
    if args.shard_data and hvd.size() > 1:
        common["shard"] = (hvd.rank(), hvd.size())
    val_common = dict(common)
    val_common.pop("shard", None)
    if args.dataset_type == "coco":
        from ..preprocessing.coco import CocoGenerator
        train_generator = CocoGenerator(args.coco_path, "train2017", transform_generator=transform_generator, **common)
        validation_generator = CocoGenerator(args.coco_path, "val2017", **val_common) \
            if os.path.exists(os.path.join(args.coco_path, "annotations", "instances_val2017.json")) else None
    elif args.dataset_type == "pascal":
        from ..preprocessing.pascal_voc import PascalVocGenerator
        train_generator = PascalVocGenerator(args.pascal_path, "trainval", transform_generator=transform_generator, **common)
        validation_generator = PascalVocGenerator(args.pascal_path, "test", **val_common)
    elif args.dataset_type == "csv":
        from ..preprocessing.csv_generator import CSVGenerator
        train_generator = CSVGenerator(args.annotations, args.classes, transform_generator=transform_generator, **common)
        validation_generator = CSVGenerator(args.val_annotations, args.classes, **val_common) if args.val_annotations else None
    elif args.dataset_type == "oid":
        from ..preprocessing.open_images import OpenImagesGenerator
        oid = dict(version=args.version, labels_filter=args.labels_filter, annotation_cache_dir=args.annotation_cache_dir,
                   fixed_labels=args.fixed_labels)
        train_generator = OpenImagesGenerator(args.main_dir, subset="train", transform_generator=transform_generator, **oid, **common)
        validation_generator = OpenImagesGenerator(args.main_dir, subset="validation", **oid, **val_common)
    elif args.dataset_type == "kitti":
        from ..preprocessing.kitti import KittiGenerator
        train_generator = KittiGenerator(args.kitti_path, subset="train", transform_generator=transform_generator, **common)
        validation_generator = KittiGenerator(args.kitti_path, subset="val", **val_common)
    elif args.dataset_type == "synthetic":
        from ..preprocessing.synthetic import SyntheticGenerator
        hw = (args.image_min_side, min(args.image_max_side, args.synthetic_width or args.image_min_side))
        train_generator = SyntheticGenerator(num_images=args.synthetic_images, image_hw=hw, classes=args.synthetic_classes,
                                             transform_generator=transform_generator, **common)
        validation_generator = SyntheticGenerator(num_images=max(2, args.synthetic_images // 4), image_hw=hw,
                                                  classes=args.synthetic_classes, seed=7, **val_common)
    else:
        raise ValueError("Invalid data type received: {}".format(args.dataset_type))
    return train_generator, validation_generator


def check_args(parsed_args):
    """Contradictions are rejected with the reference's own messages (train.py:300-326)."""
    if parsed_args.multi_gpu > 1 and parsed_args.batch_size < parsed_args.multi_gpu:
        raise ValueError("Batch size ({}) must be equal to or higher than the number of GPUs ({})".format(
            parsed_args.batch_size, parsed_args.multi_gpu))
    if parsed_args.multi_gpu > 1 and parsed_args.snapshot:
        raise ValueError("Multi GPU training ({}) and resuming from snapshots ({}) is not supported.".format(
            parsed_args.multi_gpu, parsed_args.snapshot))
    if parsed_args.multi_gpu > 1 and not parsed_args.multi_gpu_force:
        raise ValueError("Multi-GPU support is experimental, use at own risk! Run with --multi-gpu-force if you wish to continue.")
    if parsed_args.multi_gpu > 1 and getattr(parsed_args, "cpu", False):
        raise ValueError("--multi-gpu needs GPUs; drop --cpu or use the launcher (-np N) for CPU ranks.")
    if "resnet" not in parsed_args.backbone:
        warnings.warn("Using experimental backbone {}. Only resnet50 has been properly tested.".format(parsed_args.backbone))
    return parsed_args


def parse_args(args):
    parser = argparse.ArgumentParser(description="Simple training script for training a RetinaNet network.")
    subparsers = parser.add_subparsers(help="Arguments for specific dataset types.", dest="dataset_type")
    subparsers.required = True
    coco_parser = subparsers.add_parser("coco")
    coco_parser.add_argument("coco_path", help="Path to dataset directory (ie. /tmp/COCO).")
    pascal_parser = subparsers.add_parser("pascal")
    pascal_parser.add_argument("pascal_path", help="Path to dataset directory (ie. /tmp/VOCdevkit).")
    kitti_parser = subparsers.add_parser("kitti")
    kitti_parser.add_argument("kitti_path", help="Path to dataset directory (ie. /tmp/kitti).")

    def csv_list(string):
        return string.split(",")

    oid_parser = subparsers.add_parser("oid")
    oid_parser.add_argument("main_dir", help="Path to dataset directory.")
    oid_parser.add_argument("--version", help="The current dataset version is v4.", default="v4")
    oid_parser.add_argument("--labels-filter", help="A list of labels to filter.", type=csv_list, default=None)
    oid_parser.add_argument("--annotation-cache-dir", help="Path to store annotation cache.", default=".")
    oid_parser.add_argument("--fixed-labels", help="Use the exact specified labels.", default=False)
    csv_parser = subparsers.add_parser("csv")
    csv_parser.add_argument("annotations", help="Path to CSV file containing annotations for training.")
    csv_parser.add_argument("classes", help="Path to a CSV file containing class label mapping.")
    csv_parser.add_argument("--val-annotations", help="Path to CSV file containing annotations for validation (optional).")
    syn_parser = subparsers.add_parser("synthetic", help="(extension) in-memory COCO-shaped synthetic data")
    syn_parser.add_argument("--synthetic-images", type=int, default=8)
    syn_parser.add_argument("--synthetic-classes", type=int, default=80)
    syn_parser.add_argument("--synthetic-width", type=int, default=0)

    group = parser.add_mutually_exclusive_group()
    group.add_argument("--snapshot", help="Resume training from a snapshot. (extension) 'auto': the newest "
                       "checkpoint-NN.h5 under --snapshot-path if there is one, else start as if the flag were absent "
                       "-- a requeued job (reference cluster.json:6) then continues instead of restarting.")
    group.add_argument("--imagenet-weights", help="Initialize the model with pretrained imagenet weights. This is the default behaviour.", action="store_const", const=True, default=True)
    group.add_argument("--weights", help="Initialize the model with weights from a file.")
    group.add_argument("--no-weights", help="Don't initialize the model with any weights.", dest="imagenet_weights", action="store_const", const=False)

    parser.add_argument("--backbone", help="Backbone model used by retinanet.", default="resnet50", type=str)
    parser.add_argument("--batch-size", help="Size of the batches.", default=1, type=int)
    parser.add_argument("--gpu", help="Id of the GPU to use (as reported by nvidia-smi).")
    parser.add_argument("--multi-gpu", help="Number of GPUs to use for parallel processing.", type=int, default=0)
    parser.add_argument("--multi-gpu-force", help="Extra flag needed to enable (experimental) multi-gpu support.", action="store_true")
    parser.add_argument("--epochs", help="Number of epochs to train.", type=int, default=50)
    parser.add_argument("--steps", help="Number of steps per epoch.", type=int, default=10000)
    parser.add_argument("--snapshot-path", help="Path to store snapshots of models during training (defaults to './snapshots')", default="./snapshots")
    parser.add_argument("--tensorboard-dir", help="Log directory for Tensorboard output", default="./logs")
    parser.add_argument("--no-snapshots", help="Disable saving snapshots.", dest="snapshots", action="store_false")
    parser.add_argument("--no-evaluation", help="Disable per epoch evaluation.", dest="evaluation", action="store_false")
    parser.add_argument("--freeze-backbone", help="Freeze training of backbone layers.", action="store_true")
    parser.add_argument("--random-transform", help="Randomly transform image and annotations.", action="store_true")
    parser.add_argument("--image-min-side", help="Rescale the image so the smallest side is min_side.", type=int, default=800)
    parser.add_argument("--image-max-side", help="Rescale the image if the largest side is larger than max_side.", type=int, default=1333)
    # ---- extensions (not in the reference) ----
    parser.add_argument("--cpu", help="(extension) force the CPU spec engine.", action="store_true")
    parser.add_argument("--workers", help="(extension) background threads producing batches (Keras fit_generator "
                        "uses 1); 0 loads in the training thread.", type=int, default=min(8, os.cpu_count() or 1))
    parser.add_argument("--host-augment", dest="gpu_augment", action="store_false",
                        help="(extension) warp / resize images on the host with OpenCV like the reference; by default, "
                             "on a GPU, workers only decode and the warp + resize run as CUDA kernels with the same "
                             "fixed-point arithmetic.")
    parser.add_argument("--max-queue-size", help="(extension) batches kept ready ahead of the training step "
                        "(Keras default 10).", type=int, default=10)
    parser.add_argument("--engine", help="(extension) auto | native | spec", default="auto")
    parser.add_argument("--lr", help="(extension) Adam learning rate (reference hard-codes 1e-5).", type=float, default=1e-5)
    parser.add_argument("--clipnorm", help="(extension) global-norm clip (reference hard-codes 0.001).", type=float, default=0.001)
    parser.add_argument("--shard-data", help="(extension) give every rank a disjoint slice of the dataset.", action="store_true")
    parser.add_argument("--timeline", help="(extension) write a Chrome-trace timeline of host collectives (rank 0).")
    return check_args(parser.parse_args(args))


def fit_generator(trainer, generator, steps_per_epoch, epochs, callbacks, initial_epoch=0, verbose=1, workers=1,
                  max_queue_size=10, gpu_augment=False):
    """Keras ``fit_generator`` over the compact batches (reference train.py:444-450): batches are produced ahead of
    the training step by ``workers`` background threads into a queue of ``max_queue_size`` (Keras defaults 1 / 10)."""
    from ..preprocessing.prefetch import BatchPrefetcher
    params = {"epochs": epochs, "steps": steps_per_epoch, "verbose": verbose}
    cbs = list(callbacks)
    if verbose and hvd.rank() == 0:
        cbs.insert(1, ProgbarLogger(every=max(1, steps_per_epoch // 10)))
    cl = CallbackList(cbs, trainer, params)
    history = []
    cl.on_train_begin({})
    pin = trainer.device.type == "cuda"
    assembler = None
    if gpu_augment and pin and hasattr(generator, "next_raw"):
        from ..preprocessing.gpu_batch import GpuBatchAssembler
        assembler = GpuBatchAssembler(trainer.device, pad_bgr=trainer._pad_bgr() if hasattr(trainer, "_pad_bgr") else (104, 117, 124))
    batches = BatchPrefetcher(generator, workers=workers, max_queue_size=max_queue_size, pin=pin, assembler=assembler)
    for epoch in range(initial_epoch, epochs):
        cl.on_epoch_begin(epoch, {})
        sums = torch.zeros(3, device=trainer.device)
        for step in range(steps_per_epoch):
            batch = batches.next()
            out = trainer.train_step(*batch)
            sums += out.detach().to(sums.device)
            if verbose and ((step + 1) % max(1, steps_per_epoch // 10) == 0 or step + 1 == steps_per_epoch):
                m = (sums / (step + 1)).tolist()
                cl.on_batch_end(step, {"loss": m[0], "regression_loss": m[1], "classification_loss": m[2]})
        m = (sums / steps_per_epoch).tolist()
        if hasattr(trainer, "check_comm"):
            trainer.check_comm()        # a peer that vanished shows up as a timed-out flag barrier, not as a hang
        logs = {"loss": m[0], "regression_loss": m[1], "classification_loss": m[2]}
        cl.on_epoch_end(epoch, logs)
        history.append(logs)
    batches.close()
    cl.on_train_end({})
    return history


def main(args=None):
    if args is None:
        args = sys.argv[1:]
    args = parse_args(args)
    hvd.init()
    if args.timeline:
        hvd.timeline_start(args.timeline)
    backbone = models.backbone(args.backbone)
    device = get_device(args)
    check_versions(device)
    train_generator, validation_generator = create_generators(args)
    initial_epoch = 0
    state = None
    if args.snapshot == "auto":
        args.snapshot = latest_snapshot(args.snapshot_path)
        if hvd.rank() == 0:
            print("--snapshot auto: " + (f"resuming from {args.snapshot}" if args.snapshot else "no checkpoint yet, starting fresh"))
    if args.snapshot is not None:
        print("Loading model, this may take a second...")
        model, state = models.load_model(args.snapshot, backbone_name=args.backbone)
        prediction_model = models.retinanet_bbox(model=model)
        initial_epoch = int(state.get("epoch") or 0)  # the reference restarts at 0
    else:
        weights = args.weights
        if weights is None and args.imagenet_weights:
            weights = backbone.download_imagenet()
        print("Creating model, this may take a second...")
        model, _, prediction_model = create_models(backbone_retinanet=backbone.retinanet,
                                                   num_classes=train_generator.num_classes(), weights=weights,
                                                   multi_gpu=args.multi_gpu, freeze_backbone=args.freeze_backbone)
    if hvd.rank() == 0:
        print(model.summary())
    # Unlike the reference (which reloads a snapshot with a plain, un-distributed optimizer -- SURVEY
    # App. D item 1) the optimizer is always re-wrapped: create_trainer builds the distributed one.
    h, w = args.image_min_side, args.image_max_side
    if args.multi_gpu > 1:
        # single-process towers (reference train.py:86-89): N native engines in one process, P2P allreduce
        if hvd.size() > 1:
            raise ValueError("--multi-gpu towers and multi-process ranks cannot be combined (each rank owns one GPU, "
                             "reference train.py:71).")
        from ..engine.multi_tower import MultiTowerTrainer
        trainer = MultiTowerTrainer(model, args.multi_gpu, lr=args.lr, clipnorm=args.clipnorm, batch=args.batch_size,
                                    image_hw=(h, w), freeze_backbone=args.freeze_backbone)
    else:
        trainer = create_trainer(model, device, lr=args.lr, clipnorm=args.clipnorm, batch=args.batch_size,
                                 image_hw=(h, w), freeze_backbone=args.freeze_backbone, engine=args.engine)
    if state is not None:
        trainer.load_optimizer_state(state)
    callbacks = create_callbacks(model, model, prediction_model, validation_generator, args)
    history = fit_generator(trainer, train_generator, steps_per_epoch=args.steps, epochs=args.epochs, callbacks=callbacks,
                            initial_epoch=initial_epoch, verbose=1, workers=args.workers,
                            max_queue_size=args.max_queue_size, gpu_augment=args.gpu_augment)
    if args.timeline:
        hvd.timeline_stop()
    return history


if __name__ == "__main__":
    main()
