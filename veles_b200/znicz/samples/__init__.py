"""Sample workflows runnable with ``python -m veles_b200 veles_b200/znicz/samples/<name>.py -``
(stand-ins for the Znicz sample zoo — MnistSimple, CIFAR caffe, AlexNet — on synthetic data)."""
