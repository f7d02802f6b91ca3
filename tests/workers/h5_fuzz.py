"""Feeds damaged copies of the libhdf5-layout fixture to the HDF5 reader: truncations at random lengths and random byte
flips in the metadata region.  Every case must end in a clean result or a Python exception -- never a crash of the
process (this file is run as a subprocess so that a segfault fails the test instead of killing pytest)."""
import os
import random
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from batchai_retinanet_horovod_coco_b200.utils.h5 import H5File  # noqa: E402

SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fixtures", "keras_resnet50_stem_libhdf5.h5")


def walk_everything(path):
    with H5File(path, "r") as f:
        def walk(g):
            for k in f.keys(g):
                q = g.rstrip("/") + "/" + k
                for a in f.attr_names(q):
                    f.attr(q, a)
                if f.is_group(q):
                    walk(q)
                else:
                    f.dataset(q)
        for a in f.attr_names("/"):
            f.attr("/", a)
        walk("/")


def main(cases: int):
    src = open(SRC, "rb").read()
    clean = raised = 0
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "damaged.h5")
        for seed in range(cases):
            rng = random.Random(seed)
            data = bytearray(src)
            if seed % 2 == 0:
                data = data[:rng.randrange(0, len(data))]
            else:
                for _ in range(rng.randrange(1, 8)):          # the metadata lives in the first ~12 KB and in the tail
                    pos = rng.randrange(0, 12000) if rng.random() < 0.8 else rng.randrange(len(data) - 6000, len(data))
                    data[pos] = rng.randrange(256)
            with open(p, "wb") as fh:
                fh.write(bytes(data))
            try:
                walk_everything(p)
                clean += 1
            except (OSError, KeyError, ValueError, UnicodeDecodeError) as e:
                raised += 1
    print(f"FUZZ_OK clean={clean} raised={raised}")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 300)
