"""Host-side graph helpers (reference: torchpruner/utils/graph.py).

Pure Python over ``nn.Module`` trees — nothing here touches the GPU.  Behaviour, printed lines
and corner cases follow the reference so that scores are taken at the same tensors.
"""
import logging

import torch.nn as nn
from torch.nn.modules.activation import LeakyReLU, ReLU, ReLU6, RReLU, Sigmoid, Softplus, Tanh
from torch.nn.modules.batchnorm import _BatchNorm

# Same set as reference graph.py:6.
ACTIVATIONS = [ReLU, ReLU6, RReLU, LeakyReLU, Sigmoid, Softplus, Tanh]
_ACTIVATION_TYPES = tuple(ACTIVATIONS)


def find_best_module_for_attributions(model, module):
    """Module after which attributions of ``module`` are best taken (graph.py:9-34).

    Walks ``model.modules()`` in *declaration* order starting right after ``module`` and keeps
    shifting past BatchNorm / activation modules; the first module of any other kind ends the
    walk.  Two reference corner cases are kept on purpose:

    * if the walk runs off the end of the model without meeting another kind of module, the
      original ``module`` is returned (graph.py:34), not the last BatchNorm/activation seen;
    * a module that is not part of ``model`` logs an error and is returned unchanged.
    """
    ordered = list(model.modules())
    if not any(m is module for m in ordered):
        # list.index() in the reference compares with ==, which is identity for modules
        logging.error("Provided module is not in model")
        return module
    position = next(i for i, m in enumerate(ordered) if m is module)
    best = module
    for follower in ordered[position + 1:]:
        if isinstance(follower, _BatchNorm):
            print(f"BatchNorm detected: shifting evaluation after {follower}")
            best = follower
        elif isinstance(follower, _ACTIVATION_TYPES):
            print(f"Activation detected: shifting evaluation after {follower}")
            best = follower
        else:
            return best
    return module


def get_vgg_pruning_graph(vgg):
    """``[(module, [cascading modules])]`` for a VGG-style model, last layer first
    (graph.py:37-60).

    Every Linear/Conv2d opens an entry; BatchNorm2d/Dropout modules that follow it and the next
    Linear/Conv2d are its cascade.  When the next layer is met the cascade list is reversed, so
    the consumer layer comes first and BatchNorm/Dropout follow in reverse declaration order.
    The final (output) layer is dropped: pruning its outputs would change the class count.
    """
    entries = []
    for m in vgg.modules():
        if isinstance(m, (nn.Linear, nn.Conv2d)):
            if entries:
                cascade = entries[-1][1]
                cascade.append(m)
                cascade.reverse()
            entries.append((m, []))
        elif isinstance(m, (nn.BatchNorm2d, nn.Dropout)) and entries:
            entries[-1][1].append(m)
    entries.reverse()
    return entries[1:]


def get_resnet_pruning_graph(resnet):
    """``[(module, [cascading modules])]`` for a torchvision-style residual network, last block
    first.  Not in the reference, whose only graph builder is the VGG one (graph.py:37-60).

    What can be pruned without touching the residual stream are the layers INSIDE a block: in a
    bottleneck (``conv1, bn1, conv2, bn2, conv3, bn3``) the outputs of ``conv1`` feed only
    ``bn1 -> conv2`` and those of ``conv2`` only ``bn2 -> conv3``; in a basic block
    (``conv1, bn1, conv2, bn2``) the outputs of ``conv1`` feed ``bn1 -> conv2``.  The stem
    convolution (``conv1, bn1`` of the model itself) is included when the first block re-projects
    its input (``downsample``), so that its channels are read by convolutions only.  ``conv3`` /
    ``downsample`` outputs are summed into the stream every later block of the stage reads and
    writes: removing channels there would need the same indices in all of them and is left out.
    Cascade lists follow the VGG builder's order: the consumer convolution(s) first, then the
    BatchNorm."""
    entries = []
    blocks = [m for m in resnet.modules()
              if all(isinstance(getattr(m, n, None), nn.Conv2d) for n in ("conv1", "conv2"))
              and all(isinstance(getattr(m, n, None), nn.BatchNorm2d) for n in ("bn1", "bn2"))
              and m is not resnet]
    stem = getattr(resnet, "conv1", None)
    if isinstance(stem, nn.Conv2d) and isinstance(getattr(resnet, "bn1", None), nn.BatchNorm2d) \
            and blocks and getattr(blocks[0], "downsample", None) is not None:
        first = blocks[0]
        readers = [first.conv1] + [m for m in first.downsample.modules() if isinstance(m, nn.Conv2d)]
        entries.append((stem, readers + [resnet.bn1]))
    for blk in blocks:
        entries.append((blk.conv1, [blk.conv2, blk.bn1]))
        if isinstance(getattr(blk, "conv3", None), nn.Conv2d):
            entries.append((blk.conv2, [blk.conv3, blk.bn2]))
    entries.reverse()
    return entries
