"""ORACLE (test infrastructure, not product code; only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline may import it).

CPU restatement of what the reference does with the logits after the network:

    run_eval_tail      restates  Trainer.run_eval, reference train.py:337-360 (numpy: exp / sum softmax WITHOUT max
                                 subtraction, per-video mean over samples_per_video, argmax, accuracy)
    SparseModelRef     restates  SparseModel, reference train_sparse.py:89-105 (num_class nn.Linear(num_part, 1))
    sparse_test_tail   restates  SparseFusionTrainer.test, reference train_sparse.py:214-232 (mean per video, argmax)

train_sparse.py cannot be imported here (`import turtle` at :4 needs tkinter; it also imports the dataset chain), so
these are restatements read off the file; the reference has no tests for them: parity unpinned by reference tests.
"""
import numpy as np
import torch
import torch.nn as nn


def run_eval_tail(pred_score_arr: np.ndarray, true_arr: np.ndarray, samples_per_video, softmax=True):
    pred_score_arr = pred_score_arr.astype(np.float32)
    if softmax:
        pred_score_arr = np.exp(pred_score_arr) / np.sum(np.exp(pred_score_arr), axis=1, keepdims=True)
    correct_list, means, tops = [], [], []
    read_index = 0
    for num_samples in samples_per_video:
        v_begin, v_end = read_index, read_index + num_samples
        read_index += num_samples
        preds = np.mean(pred_score_arr[v_begin:v_end], axis=0)
        trues = true_arr[v_begin:v_end]
        assert np.all(trues == trues[0])
        pred_1 = np.argmax(preds, axis=0)
        correct_list.append(pred_1 == trues[0])
        means.append(preds)
        tops.append(pred_1)
    c = np.array(correct_list)
    return {"ps": pred_score_arr, "t": true_arr, "acc": c.sum() / len(c), "sv": list(samples_per_video),
            "video_mean": np.stack(means), "video_top1": np.array(tops)}


class SparseModelRef(nn.Module):
    def __init__(self, num_class, num_part):
        super().__init__()
        self.num_class = num_class
        self.num_part = num_part
        self.fcs = nn.ModuleList([nn.Linear(num_part, 1) for _ in range(num_class)])

    def forward(self, x):
        assert x.size()[1:3] == (self.num_part, self.num_class)
        y = []
        for c in range(self.num_class):
            y.append(self.fcs[c](x[:, :, c]))
        return torch.concat(y, dim=-1)


def sparse_test_tail(fused: np.ndarray, true_arr: np.ndarray, samples_per_video):
    return run_eval_tail(fused, true_arr, samples_per_video, softmax=False)
