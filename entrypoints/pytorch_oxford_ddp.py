"""Oxford-IIIT Pet DDP trainer - same CLI as the reference's 2_training_oxford-pet_ddp/pytorch_oxford_ddp.py.
    torchrun --nproc-per-node 8 entrypoints/pytorch_oxford_ddp.py --model_name swin_b --num-classes 37 \
        --height 224 --width 224 --num-epochs 15 --batch-size 80 --lr 1e-4 --backend native --bf16 --synthetic"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dist.engine.oxford import main

if __name__ == "__main__":
    main()
