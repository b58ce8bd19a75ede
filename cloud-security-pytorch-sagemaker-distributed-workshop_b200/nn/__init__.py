from .modules import (BatchNorm2d, Conv2d, CrossEntropyLoss, GlobalAvgPool, Linear, MaxPool2d, ReLU,
                      defer_bn_counters, flush_bn_counters)

__all__ = ["BatchNorm2d", "Conv2d", "CrossEntropyLoss", "GlobalAvgPool", "Linear", "MaxPool2d", "ReLU",
           "defer_bn_counters", "flush_bn_counters"]
