"""Training callbacks with the Keras / Horovod / keras-retinanet semantics that the reference
wires up in ``create_callbacks`` (``train.py:110-174``).

Every callback gets ``trainer`` (an object with ``model``, ``lr``, ``predict``, ``save`` ...) via
``set_trainer`` and the Keras hook names.  ``logs`` is a mutable dict shared along the list, in
list order, exactly like Keras.
"""

from __future__ import annotations

import os
import sys
import time

import numpy as np


class Callback:
    def __init__(self):
        self.trainer = None
        self.params = {}

    def set_trainer(self, trainer):
        self.trainer = trainer

    def set_params(self, params):
        self.params = params

    def on_train_begin(self, logs=None): pass
    def on_train_end(self, logs=None): pass
    def on_epoch_begin(self, epoch, logs=None): pass
    def on_epoch_end(self, epoch, logs=None): pass
    def on_batch_begin(self, batch, logs=None): pass
    def on_batch_end(self, batch, logs=None): pass


class CallbackList:
    def __init__(self, callbacks, trainer, params=None):
        self.callbacks = list(callbacks)
        for c in self.callbacks:
            c.set_trainer(trainer)
            c.set_params(params or {})

    def __getattr__(self, name):
        if not name.startswith("on_"):
            raise AttributeError(name)

        def call(*a, **k):
            for c in self.callbacks:
                getattr(c, name)(*a, **k)
        return call


class BroadcastGlobalVariablesCallback(Callback):
    """``hvd.callbacks.BroadcastGlobalVariablesCallback(root)``: once, at train begin, every rank takes
    rank ``root``'s weights, BN statistics and optimizer state (reference ``train.py:111``)."""

    def __init__(self, root_rank=0):
        super().__init__()
        self.root_rank = root_rank
        self.done = False

    def on_train_begin(self, logs=None):
        if not self.done:
            self.trainer.broadcast_state(self.root_rank)
            self.done = True


class MetricAverageCallback(Callback):
    """``hvd.callbacks.MetricAverageCallback``: average epoch metrics over ranks so that every rank
    (and therefore ReduceLROnPlateau) sees the same numbers (fixes SURVEY.md Appendix D item 2)."""

    def on_epoch_end(self, epoch, logs=None):
        from .. import parallel as hvd
        import torch
        if logs is None or not hvd.is_initialized() or hvd.size() == 1:
            return
        keys = sorted(k for k, v in logs.items() if isinstance(v, (int, float, np.floating)))
        if not keys:
            return
        t = torch.tensor([float(logs[k]) for k in keys], dtype=torch.float64)
        hvd.allreduce_(t, average=True, name=f"metric_avg.{epoch}")
        for k, v in zip(keys, t.tolist()):
            logs[k] = v


class ModelCheckpoint(Callback):
    """``keras.callbacks.ModelCheckpoint(filepath)`` with Keras defaults: full model every epoch,
    1-based ``{epoch:02d}`` (reference ``train.py:112-114``).  The directory is created
    (upstream's ``makedirs`` is dead code -- SURVEY.md Appendix D item 5)."""

    def __init__(self, filepath, verbose=0, save_weights_only=False):
        super().__init__()
        self.filepath, self.verbose, self.save_weights_only = filepath, verbose, save_weights_only

    def on_epoch_end(self, epoch, logs=None):
        path = self.filepath.format(epoch=epoch + 1, **(logs or {}))
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        self.trainer.save(path, epoch=epoch + 1, weights_only=self.save_weights_only)
        if self.verbose:
            print(f"\nEpoch {epoch + 1:05d}: saving model to {path}")


class TensorBoard(Callback):
    """``keras.callbacks.TensorBoard(log_dir, histogram_freq=0, write_graph=True ...)`` reduced to what the
    reference uses: epoch scalars (``loss``, ``regression_loss``, ``classification_loss``, ``lr``) and any
    metric later callbacks add (mAP / COCO stats) (reference ``train.py:119-131``)."""

    def __init__(self, log_dir="./logs", histogram_freq=0, batch_size=1, write_graph=True, write_grads=False,
                 write_images=False, embeddings_freq=0, embeddings_layer_names=None, embeddings_metadata=None):
        super().__init__()
        self.log_dir = log_dir
        self.writer = None
        self.write_graph = write_graph

    def on_train_begin(self, logs=None):
        from torch.utils.tensorboard import SummaryWriter
        os.makedirs(self.log_dir, exist_ok=True)
        self.writer = SummaryWriter(self.log_dir)
        if self.write_graph and self.trainer is not None and getattr(self.trainer, "model", None) is not None:
            self.writer.add_text("model_summary", "```\n" + self.trainer.model.summary() + "\n```", 0)

    def add_scalar(self, tag, value, step):
        if self.writer is not None:
            self.writer.add_scalar(tag, float(value), step)

    def on_epoch_end(self, epoch, logs=None):
        for k, v in (logs or {}).items():
            if isinstance(v, (int, float, np.floating)):
                self.add_scalar(k, v, epoch)
        if self.writer is not None:
            self.writer.flush()

    def on_train_end(self, logs=None):
        if self.writer is not None:
            self.writer.close()


class ReduceLROnPlateau(Callback):
    """``keras.callbacks.ReduceLROnPlateau`` (reference ``train.py:163-172``: monitor ``loss``, factor
    0.1, patience 2, epsilon 1e-4, cooldown 0, min_lr 0, mode auto)."""

    def __init__(self, monitor="loss", factor=0.1, patience=2, verbose=0, mode="auto", epsilon=1e-4, cooldown=0,
                 min_lr=0):
        super().__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0.")
        self.monitor, self.factor, self.patience, self.verbose = monitor, factor, patience, verbose
        self.epsilon, self.cooldown, self.min_lr = epsilon, cooldown, min_lr
        self.mode = mode
        self._reset()

    def _reset(self):
        if self.mode == "max" or (self.mode == "auto" and "acc" in self.monitor):
            self.monitor_op = lambda a, b: a > b + self.epsilon
            self.best = -np.inf
        else:
            self.monitor_op = lambda a, b: a < b - self.epsilon
            self.best = np.inf
        self.cooldown_counter = 0
        self.wait = 0

    def on_train_begin(self, logs=None):
        self._reset()

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        logs["lr"] = self.trainer.lr
        current = logs.get(self.monitor)
        if current is None:
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        if self.monitor_op(current, self.best):
            self.best = current
            self.wait = 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience:
                old = float(self.trainer.lr)
                if old > self.min_lr:
                    new = max(old * self.factor, self.min_lr)
                    self.trainer.lr = new
                    if self.verbose:
                        print(f"\nEpoch {epoch + 1:05d}: ReduceLROnPlateau reducing learning rate to {new}.")
                    self.cooldown_counter = self.cooldown
                    self.wait = 0


class RedirectModel(Callback):
    """keras-retinanet ``callbacks.RedirectModel(callback, model)``: forward every hook but pin the wrapped
    callback to another model / predictor (reference ``train.py:141``: evaluate the prediction graph)."""

    def __init__(self, callback, model):
        super().__init__()
        self.callback, self.redirect_model = callback, model

    def set_trainer(self, trainer):
        self.trainer = trainer
        self.callback.set_trainer(trainer)
        self.callback.predictor = self.redirect_model

    def set_params(self, params):
        self.callback.set_params(params)

    def on_train_begin(self, logs=None): self.callback.on_train_begin(logs)
    def on_train_end(self, logs=None): self.callback.on_train_end(logs)
    def on_epoch_begin(self, epoch, logs=None): self.callback.on_epoch_begin(epoch, logs)
    def on_epoch_end(self, epoch, logs=None): self.callback.on_epoch_end(epoch, logs)
    def on_batch_begin(self, batch, logs=None): self.callback.on_batch_begin(batch, logs)
    def on_batch_end(self, batch, logs=None): self.callback.on_batch_end(batch, logs)


def _shard_indices(n):
    """Evaluation is sharded over ranks and gathered (the reference evaluates the whole set on every rank
    and races on one results file -- SURVEY.md Appendix D item 7)."""
    from .. import parallel as hvd
    if hvd.is_initialized() and hvd.size() > 1:
        return list(range(hvd.rank(), n, hvd.size())), hvd.rank(), hvd.size()
    return list(range(n)), 0, 1


class Evaluate(Callback):
    """keras-retinanet ``callbacks.Evaluate``: per-class VOC AP at IoU 0.5 after every epoch; logs ``mAP``."""

    def __init__(self, generator, iou_threshold=0.5, score_threshold=0.05, max_detections=100, tensorboard=None,
                 verbose=1):
        super().__init__()
        self.generator, self.iou_threshold, self.score_threshold = generator, iou_threshold, score_threshold
        self.max_detections, self.tensorboard, self.verbose = max_detections, tensorboard, verbose
        self.predictor = None

    def on_epoch_end(self, epoch, logs=None):
        from ..utils.eval import evaluate, mean_ap
        predictor = self.predictor or self.trainer.predict
        aps = evaluate(self.generator, predictor, self.iou_threshold, self.score_threshold, self.max_detections)
        self.mean_ap = mean_ap(aps)
        if self.verbose:
            for label, (ap, n) in aps.items():
                print(f"{n:.0f} instances of class {self.generator.label_to_name(label)} with average precision: {ap:.4f}")
            print(f"mAP: {self.mean_ap:.4f}")
        if self.tensorboard is not None:
            self.tensorboard.add_scalar("mAP", self.mean_ap, epoch)
        if logs is not None:
            logs["mAP"] = self.mean_ap


class CocoEval(Callback):
    """keras-retinanet ``callbacks.coco.CocoEval``: COCO bbox evaluation after every epoch; the 12 stats
    go to TensorBoard when a writer is passed (reference ``train.py:135-138``)."""

    TAGS = ["AP @[ IoU=0.50:0.95 | area=all | maxDets=100 ]", "AP @[ IoU=0.50 | area=all | maxDets=100 ]",
            "AP @[ IoU=0.75 | area=all | maxDets=100 ]", "AP @[ IoU=0.50:0.95 | area=small | maxDets=100 ]",
            "AP @[ IoU=0.50:0.95 | area=medium | maxDets=100 ]", "AP @[ IoU=0.50:0.95 | area=large | maxDets=100 ]",
            "AR @[ IoU=0.50:0.95 | area=all | maxDets=1 ]", "AR @[ IoU=0.50:0.95 | area=all | maxDets=10 ]",
            "AR @[ IoU=0.50:0.95 | area=all | maxDets=100 ]", "AR @[ IoU=0.50:0.95 | area=small | maxDets=100 ]",
            "AR @[ IoU=0.50:0.95 | area=medium | maxDets=100 ]", "AR @[ IoU=0.50:0.95 | area=large | maxDets=100 ]"]

    def __init__(self, generator, tensorboard=None, threshold=0.05):
        super().__init__()
        self.generator, self.tensorboard, self.threshold = generator, tensorboard, threshold
        self.predictor = None
        self.stats = None

    def on_epoch_end(self, epoch, logs=None):
        from ..utils.eval import evaluate_coco
        predictor = self.predictor or self.trainer.predict
        self.stats = evaluate_coco(self.generator, predictor, self.threshold)
        if self.stats is not None:
            if self.tensorboard is not None:
                for tag, v in zip(self.TAGS, self.stats):
                    self.tensorboard.add_scalar(tag, v, epoch)
            if logs is not None:
                logs["coco_AP"] = float(self.stats[0])


class LearningRateScheduleCallback(Callback):
    """``hvd.callbacks.LearningRateScheduleCallback``: lr = initial_lr * multiplier(epoch) in [start, end)."""

    def __init__(self, multiplier, start_epoch=0, end_epoch=None, initial_lr=None):
        super().__init__()
        self.multiplier = multiplier if callable(multiplier) else (lambda e: multiplier)
        self.start_epoch, self.end_epoch, self.initial_lr = start_epoch, end_epoch, initial_lr

    def on_epoch_begin(self, epoch, logs=None):
        if self.initial_lr is None:
            self.initial_lr = self.trainer.lr
        if epoch >= self.start_epoch and (self.end_epoch is None or epoch < self.end_epoch):
            self.trainer.lr = self.initial_lr * self.multiplier(epoch)


class LearningRateWarmupCallback(LearningRateScheduleCallback):
    """``hvd.callbacks.LearningRateWarmupCallback``: ramp lr from lr/size to lr over ``warmup_epochs``."""

    def __init__(self, warmup_epochs=5, initial_lr=None, verbose=0):
        from .. import parallel as hvd
        n = hvd.size() if hvd.is_initialized() else 1
        super().__init__(lambda e: 1.0 / n * (e * (n - 1) / max(1, warmup_epochs) + 1), 0, warmup_epochs, initial_lr)


class ProgbarLogger(Callback):
    """``verbose=1`` progress lines (rank 0 only; the reference prints on all 8 ranks -- App. D item 8)."""

    def __init__(self, every=10, stream=None):
        super().__init__()
        self.every, self.stream = every, stream or sys.stdout

    def on_epoch_begin(self, epoch, logs=None):
        self.t0 = time.time()
        self.stream.write(f"Epoch {epoch + 1}/{self.params.get('epochs', '?')}\n")

    def on_batch_end(self, batch, logs=None):
        steps = self.params.get("steps", 0)
        if (batch + 1) % self.every == 0 or batch + 1 == steps:
            l = logs or {}
            self.stream.write(f"{batch + 1:>6}/{steps} - {time.time() - self.t0:.0f}s - loss: {l.get('loss', 0):.4f} - "
                              f"regression_loss: {l.get('regression_loss', 0):.4f} - classification_loss: "
                              f"{l.get('classification_loss', 0):.4f}\n")
            self.stream.flush()
