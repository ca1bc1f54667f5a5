"""b200dist - a Blackwell (sm_100a) native distributed-training framework.

Capabilities mirror the four entry points of aws-samples/sagemaker-distributed-training-pytorch-kr
(bucketed DDP, Megatron-style TP/SP/PP GPT pre-training, ZeRO-3 LLaMA/OPT fine-tuning); see DESIGN.md.
"""
__version__ = "0.1.0"
