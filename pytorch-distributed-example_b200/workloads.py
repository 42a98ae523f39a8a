"""Shapes of the gradient sets BASELINE.json names (used by bench.py and the examples).

MNIST_CNN_SHAPES: the reference `Net` (`/root/reference/mnist/main.py:130-147`) in `parameters()`
order -- 8 tensors, 431 080 fp32 elements, 1 724 320 bytes.  resnet18_shapes(): the 62 parameter
shapes of a torchvision-style ResNet-18 (configs[4]) -- 11 689 512 elements, 46 758 048 bytes.
"""

MNIST_CNN_SHAPES = [(20, 1, 5, 5), (20,), (50, 20, 5, 5), (50,), (500, 800), (500,), (10, 500), (10,)]


def resnet18_shapes(num_classes=1000):
    out = [(64, 3, 7, 7), (64,), (64,)]                      # stem conv + bn
    width_in = 64
    for stage, width in enumerate((64, 128, 256, 512)):
        for blk in range(2):
            out += [(width, width_in, 3, 3), (width,), (width,), (width, width, 3, 3), (width,), (width,)]
            if stage > 0 and blk == 0:                       # strided block: 1x1 downsample + bn
                out += [(width, width_in, 1, 1), (width,), (width,)]
            width_in = width
    return out + [(num_classes, 512), (num_classes,)]


def numel(shape):
    n = 1
    for d in shape:
        n *= int(d)
    return n
