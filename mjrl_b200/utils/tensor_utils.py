"""The one helper of mjrl/utils/tensor_utils.py the rollout path needs (:63-86): a list of
per-step info dicts -> one dict of per-path arrays, recursing into nested dicts."""
import numpy as np


def stack_tensor_list(tensor_list):
    return np.array(tensor_list)


def stack_tensor_dict_list(tensor_dict_list):
    out = {}
    if not tensor_dict_list:
        return out
    for key, first in tensor_dict_list[0].items():
        column = [step[key] for step in tensor_dict_list]
        out[key] = stack_tensor_dict_list(column) if isinstance(first, dict) else stack_tensor_list(column)
    return out
