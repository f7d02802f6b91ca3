"""B200-native data-parallel RetinaNet training framework.

Same capabilities as msalvaris/batchai_retinanet_horovod_coco (reference ``train.py``:
Keras RetinaNet + Horovod), re-designed for Blackwell: PyTorch is the module/autograd
layer, hot ops are hand-written sm_100a CUDA kernels (``csrc/``) reached through a C ABI
(``include/``), and the gradient allreduce is a peer-to-peer NVLink kernel.

Sub-packages
------------
models        RetinaNet / ResNet spec layer (reference ``train.py:43,81-107``)
ops           anchors, losses, box coding, NMS, native kernel bindings
engine        native fused forward/backward executor (the B200 hot path)
parallel      Horovod-equivalent runtime (reference ``train.py:20-21,103,111``)
preprocessing dataset generators (reference ``train.py:177-297``)
callbacks     checkpoint / TensorBoard / eval / LR plateau (reference ``train.py:110-174``)
utils         HDF5 subset, COCO eval, transforms, timeline, config
bin           CLI entry points (reference ``train.py:300-450``)
"""

__version__ = "0.1.0"
