"""Single-process MNIST control script (reference: 1_training_mnist_ddp/pytorch_mnist.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dist.engine.mnist import main

if __name__ == "__main__":
    main(ddp=False)
