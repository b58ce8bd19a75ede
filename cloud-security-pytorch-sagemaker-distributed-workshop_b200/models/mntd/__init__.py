from .models import MODELS, AudioModel, CIFAR10Model, MNISTModel, RTNLPModel, WordEmb, mel_filterbank
from .trojan import random_troj_setting, stamp_batch, troj_gen_func

__all__ = ["MODELS", "AudioModel", "CIFAR10Model", "MNISTModel", "RTNLPModel", "WordEmb", "mel_filterbank",
           "random_troj_setting", "stamp_batch", "troj_gen_func"]
