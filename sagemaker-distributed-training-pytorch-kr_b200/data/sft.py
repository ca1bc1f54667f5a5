"""Supervised fine-tuning data for the Alpaca workload (reference: 4_training_alpaca_deepspeed/train.py - PROMPT_DICT
:58-69, _tokenize_fn :115-136, preprocess :139-151, SupervisedDataset :154-180, DataCollatorForSupervisedDataset
:183-199; `utils.jload` from stanford_alpaca, which the reference imports but does not ship).

Also provides the synthetic fixed-length instruction batches BASELINE.json config 5 asks for, and a dependency-free
whitespace/byte tokenizer so the pipeline runs on a box without tokenizer files (a HF tokenizer is used when given).
"""
from __future__ import annotations

import copy
import io
import json
import logging
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch

IGNORE_INDEX = -100
DEFAULT_PAD_TOKEN, DEFAULT_EOS_TOKEN, DEFAULT_BOS_TOKEN, DEFAULT_UNK_TOKEN = "[PAD]", "</s>", "<s>", "<unk>"
PROMPT_DICT = {
    "prompt_input": ("Below is an instruction that describes a task, paired with an input that provides further context. "
                     "Write a response that appropriately completes the request.\n\n"
                     "### Instruction:\n{instruction}\n\n### Input:\n{input}\n\n### Response:"),
    "prompt_no_input": ("Below is an instruction that describes a task. "
                        "Write a response that appropriately completes the request.\n\n"
                        "### Instruction:\n{instruction}\n\n### Response:"),
}


def jload(f, mode="r"):
    """Load a .json file into a list / dict (stanford_alpaca utils.jload)."""
    if not isinstance(f, io.IOBase):
        f = open(f, mode=mode)
    try:
        return json.load(f)
    finally:
        f.close()


class ByteTokenizer:
    """Minimal tokenizer with the attributes the pipeline touches (pad/eos/bos/unk, model_max_length, __len__,
    add_special_tokens, __call__ with truncation).  ids: 0..255 bytes, then specials."""

    def __init__(self, model_max_length: int = 512, vocab_size: int = 260):
        self.model_max_length = model_max_length
        self.special: Dict[str, int] = {}
        self.pad_token = self.eos_token = self.bos_token = self.unk_token = None
        self._base = 256
        self._min_vocab = vocab_size
        self.padding_side = "right"

    def add_special_tokens(self, d: Dict[str, str]) -> int:
        n = 0
        for k, tok in d.items():
            if tok not in self.special:
                self.special[tok] = self._base + len(self.special)
                n += 1
            setattr(self, k, tok)
        return n

    def __len__(self):
        return max(self._min_vocab, self._base + len(self.special))

    @property
    def pad_token_id(self):
        return self.special.get(self.pad_token)

    @property
    def eos_token_id(self):
        return self.special.get(self.eos_token)

    def encode(self, text: str) -> List[int]:
        ids: List[int] = []
        i = 0
        while i < len(text):
            for tok, tid in self.special.items():
                if text.startswith(tok, i):
                    ids.append(tid)
                    i += len(tok)
                    break
            else:
                ids.extend(text[i].encode("utf-8"))
                i += 1
        return ids

    def __call__(self, text: str, return_tensors="pt", padding="longest", max_length=None, truncation=True):
        ids = self.encode(text)
        if truncation and max_length:
            ids = ids[:max_length]
        return type("Enc", (), {"input_ids": torch.tensor([ids], dtype=torch.long)})()


def _tokenize_fn(strings: Sequence[str], tokenizer) -> Dict:
    toks = [tokenizer(t, return_tensors="pt", padding="longest", max_length=tokenizer.model_max_length, truncation=True) for t in strings]
    input_ids = labels = [t.input_ids[0] for t in toks]
    lens = [int(t.input_ids.ne(tokenizer.pad_token_id).sum().item()) if tokenizer.pad_token_id is not None else t.input_ids.shape[1] for t in toks]
    return dict(input_ids=input_ids, labels=labels, input_ids_lens=lens, labels_lens=lens)


def preprocess(sources: Sequence[str], targets: Sequence[str], tokenizer) -> Dict:
    """tokenize prompt+response; mask the prompt part of the labels with IGNORE_INDEX."""
    examples = [s + t for s, t in zip(sources, targets)]
    ex_tok, src_tok = _tokenize_fn(examples, tokenizer), _tokenize_fn(sources, tokenizer)
    input_ids = ex_tok["input_ids"]
    labels = copy.deepcopy(input_ids)
    for label, n in zip(labels, src_tok["input_ids_lens"]):
        label[:n] = IGNORE_INDEX
    return dict(input_ids=input_ids, labels=labels)


class SupervisedDataset(torch.utils.data.Dataset):
    """Alpaca-format JSON list of {instruction, input, output}.  Tokenisation is lazy + cached per example (the
    reference tokenises all 52k examples eagerly on every rank)."""

    def __init__(self, data_path: str, tokenizer, eager: bool = False):
        logging.warning("Loading data...")
        self.records = jload(data_path)
        self.tokenizer = tokenizer
        self._cache: Dict[int, Dict[str, torch.Tensor]] = {}
        if eager:
            logging.warning("Tokenizing inputs... This may take some time...")
            for i in range(len(self.records)):
                self[i]

    def __len__(self):
        return len(self.records)

    def __getitem__(self, i) -> Dict[str, torch.Tensor]:
        if i not in self._cache:
            ex = self.records[i]
            src = (PROMPT_DICT["prompt_input"] if ex.get("input", "") != "" else PROMPT_DICT["prompt_no_input"]).format_map(ex)
            tgt = f"{ex['output']}{self.tokenizer.eos_token}"
            d = preprocess([src], [tgt], self.tokenizer)
            if not getattr(self, "_warned_empty", False) and bool((d["labels"][0] == IGNORE_INDEX).all()):
                # the prompt alone fills model_max_length: no supervised token is left and the loss of such a batch is NaN (the
                # reference behaves the same way, silently)
                self._warned_empty = True
                logging.warning("example %d has no target tokens left after truncation to model_max_length=%s: its loss is undefined "
                                "(raise --model_max_length)", i, getattr(self.tokenizer, "model_max_length", "?"))
            self._cache[i] = dict(input_ids=d["input_ids"][0], labels=d["labels"][0])
        return self._cache[i]


class SyntheticInstructionDataset(torch.utils.data.Dataset):
    """fixed-length synthetic instruction samples: random ids, the first third masked like a prompt."""

    def __init__(self, n: int, seq_len: int, vocab_size: int, seed: int = 0):
        self.n, self.s, self.v, self.seed = n, seq_len, vocab_size, seed

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 7919 + i)
        ids = torch.randint(3, self.v, (self.s,), generator=g)
        labels = ids.clone()
        labels[: self.s // 3] = IGNORE_INDEX
        return dict(input_ids=ids, labels=labels)


@dataclass
class DataCollatorForSupervisedDataset:
    """right-pad to the batch-longest sequence; attention_mask = input_ids != pad (train.py:183-199)."""
    tokenizer: object

    def __call__(self, instances: Sequence[Dict]) -> Dict[str, torch.Tensor]:
        input_ids = [x["input_ids"] for x in instances]
        labels = [x["labels"] for x in instances]
        pad = self.tokenizer.pad_token_id if self.tokenizer.pad_token_id is not None else 0
        input_ids = torch.nn.utils.rnn.pad_sequence(input_ids, batch_first=True, padding_value=pad)
        labels = torch.nn.utils.rnn.pad_sequence(labels, batch_first=True, padding_value=IGNORE_INDEX)
        return dict(input_ids=input_ids, labels=labels, attention_mask=input_ids.ne(pad))


def smart_tokenizer_and_embedding_resize(special_tokens_dict: Dict, tokenizer, model) -> int:
    """add missing special tokens; new embedding rows are initialised to the mean of the old ones (train.py:92-112)."""
    n_new = tokenizer.add_special_tokens(special_tokens_dict)
    new_size = len(tokenizer)
    emb = model.embed_tokens
    old = emb.shape[0]
    if new_size > old:
        def grow(p):
            with torch.no_grad():
                mean = p.data[:old].float().mean(0, keepdim=True).to(p.dtype)
                return torch.nn.Parameter(torch.cat([p.data, mean.expand(new_size - old, -1)], 0))
        model.embed_tokens = grow(model.embed_tokens)
        if getattr(model, "lm_head", None) is not None:
            model.lm_head = grow(model.lm_head)
        model.config.vocab_size = new_size
    return n_new


def make_supervised_data_module(tokenizer, data_args, training_args=None, vocab_size: Optional[int] = None) -> Dict:
    if training_args is not None and (training_args.synthetic_data or not data_args.data_path):
        ds = SyntheticInstructionDataset(training_args.synthetic_samples, training_args.model_max_length, vocab_size or len(tokenizer), training_args.seed)
    else:
        ds = SupervisedDataset(data_args.data_path, tokenizer)
    return dict(train_dataset=ds, eval_dataset=None, data_collator=DataCollatorForSupervisedDataset(tokenizer))
