"""Logging in the reference's format.

The reference logs to stdout at DEBUG through a ``StreamHandler(sys.stdout)``
(``notebooks/code/cifar10-distributed-native-cpu.py:17-19``).  The exact message formats are the
user-visible surface SageMaker metric regexes key on, so they are centralised here.
"""
from __future__ import annotations

import logging
import sys

_CONFIGURED = False


def get_logger(name: str = "b200ddp") -> logging.Logger:
    global _CONFIGURED
    root = logging.getLogger("b200ddp")
    if not _CONFIGURED:
        root.setLevel(logging.DEBUG)
        h = logging.StreamHandler(sys.stdout)
        h.setFormatter(logging.Formatter("%(message)s"))
        root.addHandler(h)
        root.propagate = False           # the reference's lines appear twice; ours once
        _CONFIGURED = True
    return root if name == "b200ddp" else root.getChild(name.split(".")[-1])


def fmt_train_line(epoch: int, seen: int, shard: int, pct: float, loss: float) -> str:
    """``Train Epoch: e [seen/shard (pct%)] Loss: x`` (…native-cpu.py:157-166)."""
    return "Train Epoch: {} [{}/{} ({:.0f}%)] Loss: {:.6f}".format(epoch, seen, shard, pct, loss)


def fmt_test_line(avg_loss: float, accuracy: float) -> str:
    """``Test set: Average loss: x, Accuracy: y`` (…native-cpu.py:190-194)."""
    return "Test set: Average loss: {:.4f}, Accuracy: {:.2f}\n".format(avg_loss, accuracy)


def fmt_shard_line(n_shard: int, n_total: int, what: str) -> str:
    """``Processes a/b (p%) of train data`` (…native-cpu.py:122-136)."""
    return "Processes {}/{} ({:.0f}%) of {} data".format(n_shard, n_total, 100.0 * n_shard / max(n_total, 1), what)
