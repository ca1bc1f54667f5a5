"""MNIST DDP trainer - same CLI as the reference's 1_training_mnist_ddp/pytorch_mnist_ddp.py.
Launch:  torchrun --nproc-per-node N entrypoints/pytorch_mnist_ddp.py --epochs 5 --backend nccl
(or under mpirun: OMPI_COMM_WORLD_* are honoured)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dist.engine.mnist import main

if __name__ == "__main__":
    main(ddp=True)
