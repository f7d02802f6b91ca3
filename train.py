#!/usr/bin/env python
"""Entry point with the reference's name and CLI (``/root/reference/train.py``): forwards to
``batchai_retinanet_horovod_coco_b200.bin.train.main``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from batchai_retinanet_horovod_coco_b200.bin.train import main  # noqa: E402

if __name__ == "__main__":
    main()
