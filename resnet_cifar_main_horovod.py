#!/usr/bin/env python
"""resnet_cifar_main_horovod - same name and flags as the reference's entry point; see
distributed_resnet_tensorflow_b200/cli/__init__.py:cifar_main_horovod."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from distributed_resnet_tensorflow_b200.cli import cifar_main_horovod  # noqa: E402

if __name__ == "__main__":
    cifar_main_horovod(sys.argv[1:])
