"""Model zoo: layer configurations of the BASELINE.json configs and workflow factories.

  * ``mnist_mlp``      — MNIST-shape MLP 784-100-10 (All2AllTanh + All2AllSoftmax)
  * ``cifar_convnet``  — CIFAR-10 Znicz/Caffe-style convnet (3x conv+pool, 32x32x3)
  * ``alexnet``        — AlexNet, ImageNet shape 227x227x3, 1000 classes (~61 M parameters)
  * ``wide_mlp_ae``    — 4x4096 All2All autoencoder stack (gradient all-reduce stress test)
  * ``lstm_classifier`` — LSTM over 16 x 64 sequences + softmax (recurrent layers, BPTT)
  * ``mnist_lstm``     — MNIST-shape images as 28-step sequences of 28 pixels, LSTM + softmax
Datasets are synthetic with the named shapes (no network on the build/GPU boxes).
"""
from __future__ import annotations

from .configs import (mnist_mlp_layers, cifar_convnet_layers, alexnet_layers, wide_mlp_ae_layers,  # noqa: F401
                      lstm_classifier_layers, mnist_lstm_layers,
                      MODEL_CONFIGS, build_workflow, smoke_step)
