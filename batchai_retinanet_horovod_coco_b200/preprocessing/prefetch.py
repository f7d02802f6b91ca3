"""Background batch production for the training loop.

Keras ``fit_generator`` (reference ``train.py:444-450``) pulls batches from a queue that one background
thread fills (``workers=1, max_queue_size=10``).  A B200 consumes ~450 images/s per GPU, far beyond one Python
thread of JPEG decoding, so the queue here is fed by several workers: image decode, affine warp and resize
are OpenCV calls that release the GIL, which is why threads (not processes) are enough and the batches can be
handed over without a copy.  Each worker calls the generator's thread-safe ``next_compact`` (the group cursor
and the random-transform stream are guarded by the generator's lock), so with one worker the order is exactly
the generator's order and with several it is the order of completion, as in Keras.
"""

from __future__ import annotations

import queue
import threading


class BatchPrefetcher:
    """Iterator over ``generator.next_compact(pin=...)`` results produced ahead of time by worker threads."""

    _STOP = object()

    def __init__(self, generator, workers: int = 4, max_queue_size: int = 10, pin: bool = False, assembler=None):
        """``assembler`` (``preprocessing.gpu_batch.GpuBatchAssembler``): the workers then only DECODE
        (``generator.next_raw()``; OpenCV's decoder releases the GIL, so threads scale), upload the raw pixels and
        launch the warp / resize kernels on their own CUDA stream; the consumer only waits for an event."""
        self.generator = generator
        self.pin = pin
        self.assembler = assembler
        self.workers = max(0, int(workers))
        self._q: "queue.Queue" = queue.Queue(maxsize=max(1, int(max_queue_size)))
        self._stop = threading.Event()
        self._threads = []
        for i in range(self.workers):
            t = threading.Thread(target=self._work, name=f"batch-prefetch-{i}", daemon=True)
            t.start()
            self._threads.append(t)

    def _work(self):
        while not self._stop.is_set():
            try:
                if self.assembler is not None:
                    item = self.assembler.assemble_async(self.generator.next_raw())
                else:
                    item = self.generator.next_compact(pin=self.pin)
            except BaseException as e:  # noqa: BLE001 - handed to the consumer, which re-raises it
                item = e
            while not self._stop.is_set():
                try:
                    self._q.put(item, timeout=0.1)
                    break
                except queue.Full:
                    continue
            if isinstance(item, BaseException):
                return

    def next(self):
        if self.workers == 0:                  # synchronous mode: the caller's thread does the work
            if self.assembler is not None:
                return self.assembler.assemble(self.generator.next_raw())
            return self.generator.next_compact(pin=self.pin)
        item = self._q.get()
        if isinstance(item, BaseException):
            self.close()
            raise item
        return self.assembler.finish(item) if self.assembler is not None else item

    __next__ = next

    def __iter__(self):
        return self

    def close(self):
        self._stop.set()
        try:
            while True:
                self._q.get_nowait()
        except queue.Empty:
            pass
        for t in self._threads:
            t.join(timeout=5)
        self._threads = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
