"""Host helpers of the path from Utils/Utils.py."""
import numpy as np


def split_func(x, split_pos):
    """Utils/Utils.py:28-35: undo the per-device packing along the time axis."""
    rst = []
    start = 0
    for i in range(split_pos.shape[0]):
        rst.append(x[:, start:start + split_pos[i]])
        start += split_pos[i]
    return rst


def shape_list(x):
    return list(np.shape(x))
