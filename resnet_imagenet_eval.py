#!/usr/bin/env python
"""resnet_imagenet_eval - same name and flags as the reference's entry point; see
distributed_resnet_tensorflow_b200/cli/__init__.py:imagenet_eval."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from distributed_resnet_tensorflow_b200.cli import imagenet_eval  # noqa: E402

if __name__ == "__main__":
    imagenet_eval(sys.argv[1:])
