"""Mirror of resnet/models.py's hot functions (the flow encoder body).

The reference builds Keras layers with fresh weights; here the blocks live inside a
FlowNet (weights in its flat store), so the functions take ``net=``."""
from __future__ import annotations


def resnet_block_simple(input_tensor, downsample, layer_idx, block_idx, net=None):
    """resnet/models.py:8-36."""
    assert net is not None, "pass net= (FlowNet owning the weights)"
    blk = net.encoder.blocks[(layer_idx - 2) * 2 + (block_idx - 1)]
    assert blk.down == bool(downsample)
    return blk.forward(input_tensor)


def resnet_layer_simple(input_tensor, num_blocks, downsample, layer_idx, net=None):
    """resnet/models.py:75-80."""
    x = input_tensor
    for i in range(1, num_blocks + 1):
        x = resnet_block_simple(x, i == 1 and downsample, layer_idx, i, net=net)
    return x
