"""The four MNTD task models with the reference's layer names (so the 111 shipped MNIST checkpoints and
any ``.model`` file written by the reference load unchanged) — ``notebooks/code/model_lib/``:
``mnist_cnn_model.py:6-35``, ``cifar10_cnn_model.py:6-41``, ``audio_rnn_model.py:7-45``,
``rtNLP_cnn_model.py:6-70``.

Interface kept: ``Model(gpu=False)`` with ``forward`` / ``loss`` (and ``emb_forward`` / ``emb_info`` for the
NLP model).  Convolutions and linears are ``b200ddp.nn`` modules: on a CUDA input they run the tensor-core
kernels (channels padded to 8), on CPU the reference fp32 math.  ``gpu=True`` moves the model to the
current CUDA device; inputs are moved to the model's device in ``forward`` as the reference does.

Bit-rot fixed: ``torch.stft(..., return_complex=True)`` and a built-in mel filterbank
(librosa is not installed here and ``librosa.filters.mel`` positional arguments no longer exist upstream).
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ... import nn as bnn
from ...ops import functional as BF


def _dev(module: nn.Module) -> torch.device:
    return next(module.parameters()).device


class _Base(nn.Module):
    def __init__(self, gpu: bool = False):
        super().__init__()
        self.gpu = gpu

    def _finish(self):
        if self.gpu:
            self.cuda()

    def _in(self, x):
        return x.to(_dev(self)) if x.device != _dev(self) else x


class MNISTModel(_Base):
    """2 x conv5x5 + 2 x linear (mnist_cnn_model.py:11-15,25-28)."""
    input_size = (1, 28, 28)
    class_num = 10

    def __init__(self, gpu=False):
        super().__init__(gpu)
        self.conv1 = bnn.Conv2d(1, 16, kernel_size=5, padding=0, relu=True)
        self.conv2 = bnn.Conv2d(16, 32, kernel_size=5, padding=0, relu=True)
        self.max_pool = bnn.MaxPool2d(kernel_size=2, stride=2)
        self.fc = bnn.Linear(32 * 4 * 4, 512, relu=True, nhwc_from=(32, 4, 4))
        self.output = bnn.Linear(512, 10)
        self._finish()

    def forward(self, x):
        x = self._in(x)
        x = self.max_pool(self.conv1(x))
        x = self.max_pool(self.conv2(x))
        if not x.is_cuda:
            x = x.reshape(x.shape[0], 32 * 4 * 4)
        x = self.output(self.fc(x))
        return x[:, :10].float() if x.is_cuda else x

    def loss(self, pred, label):
        return BF.cross_entropy(pred.contiguous(), label.to(pred.device)) if pred.is_cuda else F.cross_entropy(pred, label)


class CIFAR10Model(_Base):
    """4 x conv3x3 + 3 x linear, dropout 0.5 (cifar10_cnn_model.py:11-18,23-36)."""
    input_size = (3, 32, 32)
    class_num = 10

    def __init__(self, gpu=False):
        super().__init__(gpu)
        self.conv1 = bnn.Conv2d(3, 32, kernel_size=3, padding=1, relu=True)
        self.conv2 = bnn.Conv2d(32, 32, kernel_size=3, padding=1, relu=True)
        self.conv3 = bnn.Conv2d(32, 64, kernel_size=3, padding=1, relu=True)
        self.conv4 = bnn.Conv2d(64, 64, kernel_size=3, padding=1, relu=True)
        self.max_pool = bnn.MaxPool2d(kernel_size=2, stride=2)
        self.linear = bnn.Linear(64 * 8 * 8, 256, relu=True, nhwc_from=(64, 8, 8))
        self.fc = bnn.Linear(256, 256, relu=True)
        self.output = bnn.Linear(256, 10)
        self._finish()

    def forward(self, x):
        x = self._in(x)
        x = self.conv1(x)
        x = self.max_pool(self.conv2(x))
        x = self.conv3(x)
        x = self.max_pool(self.conv4(x))
        if not x.is_cuda:
            x = x.reshape(x.shape[0], 64 * 8 * 8)
        x = self.linear(x)
        x = F.dropout(self.fc(x), 0.5, training=self.training)
        x = self.output(x)
        return x[:, :10].float() if x.is_cuda else x

    def loss(self, pred, label):
        return BF.cross_entropy(pred.contiguous(), label.to(pred.device)) if pred.is_cuda else F.cross_entropy(pred, label)


def mel_filterbank(sr: int = 16000, n_fft: int = 2048, n_mels: int = 40) -> torch.Tensor:
    """Slaney-style mel filterbank (what ``librosa.filters.mel(16000, 2048, n_mels=40)`` returns)."""
    def hz_to_mel(f):
        f = np.asarray(f, dtype=np.float64)
        f_sp = 200.0 / 3
        mels = f / f_sp
        min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
        min_log_mel = min_log_hz / f_sp
        return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, mels)

    def mel_to_hz(m):
        m = np.asarray(m, dtype=np.float64)
        f_sp = 200.0 / 3
        min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
        min_log_mel = min_log_hz / f_sp
        return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)

    fftfreqs = np.linspace(0, sr / 2, 1 + n_fft // 2)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(0.0), hz_to_mel(sr / 2.0), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    weights = np.zeros((n_mels, 1 + n_fft // 2))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2: n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, None]
    return torch.from_numpy(weights.astype(np.float32))


class AudioModel(_Base):
    """STFT -> 40 mel bands (dB) -> 2-layer LSTM(100) -> attention pooling -> linear(10)
    (audio_rnn_model.py:11-13,25-39)."""
    input_size = (16000,)
    class_num = 10

    def __init__(self, gpu=False):
        super().__init__(gpu)
        self.lstm = nn.LSTM(input_size=40, hidden_size=100, num_layers=2, batch_first=True)
        self.lstm_att = nn.Linear(100, 1)
        self.output = nn.Linear(100, 10)
        self.register_buffer("_window", torch.hann_window(2048), persistent=False)
        self.register_buffer("_mel", mel_filterbank(16000, 2048, 40), persistent=False)
        self._finish()

    def features(self, x):
        spec = torch.stft(x, n_fft=2048, window=self._window, return_complex=True)
        power = spec.real ** 2 + spec.imag ** 2                       # == norm(p=2, dim=-1) ** 2 of the old API
        mel_f = torch.matmul(self._mel, power)
        mel_feature = 10 * torch.log10(torch.clamp(mel_f, min=1e-10))
        return (mel_feature.transpose(-1, -2) + 50) / 50

    def forward(self, x):
        x = self._in(x)
        if x.is_cuda:
            # native path: FFT + mel + dB in one kernel (with its backward to the waveform, which the
            # meta-classifier's query tuning needs), LSTM recurrence with W_hh in registers
            feats = BF.mel_features(x, self._window, self._mel)
            lstm_out = BF.lstm(feats, self.lstm)
            att = BF.linear_f32(lstm_out, self.lstm_att.weight, self.lstm_att.bias).squeeze(2)
            emb = (lstm_out * F.softmax(att, dim=1).unsqueeze(2)).sum(1)
            return BF.linear_f32(emb, self.output.weight, self.output.bias)
        lstm_out, _ = self.lstm(self.features(x))
        att_val = F.softmax(self.lstm_att(lstm_out).squeeze(2), dim=1)
        emb = (lstm_out * att_val.unsqueeze(2)).sum(1)
        return self.output(emb)

    def loss(self, pred, label):
        return F.cross_entropy(pred, label.to(pred.device))


class WordEmb:
    """Frozen word2vec table; deliberately not an ``nn.Module`` so it is neither saved nor trained
    (rtNLP_cnn_model.py:6-19).  Falls back to a seeded random table when ``saved_emb.npy`` is absent
    (the reference does not ship it)."""

    def __init__(self, gpu, emb_path, vocab=18000, dim=300):
        if emb_path and os.path.exists(emb_path):
            w2v_value = np.load(emb_path)
        else:
            w2v_value = np.random.RandomState(0).normal(0, 0.3, size=(vocab, dim)).astype(np.float32)
        self.embed = nn.Embedding(*w2v_value.shape)
        self.embed.weight.data = torch.from_numpy(np.asarray(w2v_value, dtype=np.float32))
        self.embed.weight.requires_grad_(False)
        self.gpu = gpu
        if gpu:
            self.embed.cuda()

    def calc_emb(self, x):
        return self.embed(x.to(self.embed.weight.device))


class RTNLPModel(_Base):
    """Kim-CNN over frozen embeddings: 3 convs (k x 300), max-over-time, linear(300 -> 1)
    (rtNLP_cnn_model.py:22-70)."""
    input_size = (1, 10, 300)
    class_num = 1

    def __init__(self, gpu=False, emb_path="./raw_data/rt_polarity/saved_emb.npy"):
        super().__init__(gpu)
        self.embed_static = WordEmb(gpu, emb_path=emb_path)
        self.conv1_3 = nn.Conv2d(1, 100, (3, 300))
        self.conv1_4 = nn.Conv2d(1, 100, (4, 300))
        self.conv1_5 = nn.Conv2d(1, 100, (5, 300))
        self.output = nn.Linear(3 * 100, 1)
        self._finish()

    @staticmethod
    def conv_and_pool(x, conv):
        x = F.relu(conv(x)).squeeze(3)
        return F.max_pool1d(x, x.size(2)).squeeze(2)

    def forward(self, x):
        x = self._in(x)
        table = self.embed_static.embed.weight
        if x.is_cuda and table.is_cuda:
            from ...ops import native as nv
            B, Ls = x.shape
            return self._native_head(nv.gather_rows(table, x.long(), pad_rows=4), B, Ls)      # embedding gather kernel (W20)
        return self.emb_forward(self.embed_static.calc_emb(x).unsqueeze(1))

    def _native_head(self, e_flat, B, Ls):
        """e_flat [B*Ls + 4, 300] (4 zero rows of padding): three (k x 300) convs as GEMMs over overlapping rows +
        fused bias/ReLU/max-over-time, dropout, fp32 linear."""
        x = torch.cat([BF.text_conv_pool(e_flat, B, Ls, c) for c in (self.conv1_3, self.conv1_4, self.conv1_5)], dim=1)
        x = F.dropout(x, 0.5, training=self.training)
        return BF.linear_f32(x, self.output.weight, self.output.bias).squeeze(1)

    def emb_forward(self, x):
        x = self._in(x)
        if x.is_cuda:
            B, _, Ls, D = x.shape
            e_flat = torch.cat([x.float().reshape(B * Ls, D), x.new_zeros((4, D), dtype=torch.float32)], 0)
            return self._native_head(e_flat, B, Ls)
        x = torch.cat([self.conv_and_pool(x, c) for c in (self.conv1_3, self.conv1_4, self.conv1_5)], dim=1)
        x = F.dropout(x, 0.5, training=self.training)
        return self.output(x).squeeze(1)

    def loss(self, pred, label):
        return F.binary_cross_entropy_with_logits(pred, label.to(pred.device).float())

    def emb_info(self):
        emb_matrix = self.embed_static.embed.weight.data
        return emb_matrix.mean(0), emb_matrix.std(0, unbiased=True)


MODELS = {"mnist": MNISTModel, "cifar10": CIFAR10Model, "audio": AudioModel, "rtNLP": RTNLPModel}
