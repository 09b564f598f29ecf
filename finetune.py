"""Pretrain / finetune GPT, Llama (v1/v2) and Falcon. Same flags as the reference's
finetune.py (:17-150): ``--model_name {gpt,llama,llama2,falcon,codellama}``,
``--model_type``, plus ``--synthetic_data`` for runs without a corpus."""
from functools import partial

import torch

from epfl_megatron_b200 import get_args, get_timers, get_tokenizer, print_rank_0
from epfl_megatron_b200.initialize import initialize_megatron
from epfl_megatron_b200.models import FalconModel, GPTModel, LlamaModel, ModelType
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.parallel.data import broadcast_data
from epfl_megatron_b200.training import pretrain
from epfl_megatron_b200.utils import average_losses_across_data_parallel_group, get_ltor_masks_and_position_ids


def model_provider(pre_process=True, post_process=True):
    print_rank_0("Building model ...")
    args = get_args()
    if args.model_name == "gpt":
        cls = GPTModel
    elif args.model_name == "falcon":
        cls = FalconModel
    elif args.model_name in {"llama", "llama2", "codellama"}:
        cls = partial(LlamaModel, version=1 if args.model_name == "llama" else 2)
    else:
        raise KeyError(f"Unknown model {args.model_name}")
    if isinstance(args.model_type, ModelType):
        model_type = args.model_type
    elif args.model_type == "encoder_or_decoder":
        model_type = ModelType.encoder_or_decoder
    elif args.model_type == "encoder_and_decoder":
        model_type = ModelType.encoder_and_decoder
    else:
        raise KeyError(f"Unsupported model_type {args.model_type}")
    return cls(num_tokentypes=0, parallel_output=True, pre_process=pre_process, post_process=post_process,
               model_type=model_type)


def get_batch(data_iterator):
    """tokens/labels shifted by one, loss mask and position ids. The [1,1,s,s] attention
    mask is only built when the unfused attention path will read it."""
    args = get_args()
    tokenizer = get_tokenizer()
    data = next(data_iterator) if data_iterator is not None else None
    data_b = broadcast_data(["text"], data, torch.int64)
    tokens_ = data_b["text"].long()
    labels = tokens_[:, 1:].contiguous()
    tokens = tokens_[:, :-1].contiguous()
    need_mask = (not args.use_flash_attn) or args.reset_attention_mask
    attention_mask, loss_mask, position_ids = get_ltor_masks_and_position_ids(
        tokens, tokenizer.eod, args.reset_position_ids, args.reset_attention_mask, args.eod_mask_loss,
        build_attention_mask=need_mask)
    return tokens, labels, loss_mask, attention_mask, position_ids


def loss_func(loss_mask, output_tensor):
    losses = output_tensor.float()
    loss_mask = loss_mask.view(-1).float()
    loss = torch.sum(losses.view(-1) * loss_mask) / loss_mask.sum()
    averaged = average_losses_across_data_parallel_group([loss])
    return loss, {"lm loss": averaged[0]}


def forward_step(data_iterator, model):
    timers = get_timers()
    timers("batch-generator", log_level=2).start()
    tokens, labels, loss_mask, attention_mask, position_ids = get_batch(data_iterator)
    timers("batch-generator").stop()
    args = get_args()
    pos = position_ids if args.reset_position_ids else None
    out = model(tokens, pos if pos is not None else position_ids, attention_mask, labels=labels)
    return out, partial(loss_func, loss_mask)


def train_valid_test_datasets_provider(train_val_test_num_samples):
    args = get_args()
    if args.synthetic_data:
        from epfl_megatron_b200.data.synthetic import build_synthetic_datasets
        return build_synthetic_datasets(args, train_val_test_num_samples)
    from epfl_megatron_b200.data.gpt_dataset import build_train_valid_test_datasets
    print_rank_0("> building train, validation, and test datasets ...")
    train_ds, valid_ds, test_ds = build_train_valid_test_datasets(
        data_prefix=args.data_path, data_impl=args.data_impl, splits_string=args.split,
        train_valid_test_num_samples=train_val_test_num_samples, seq_length=args.seq_length, seed=args.seed,
        skip_warmup=(not args.mmap_warmup), train_data_prefix=args.train_data_path,
        valid_data_prefix=args.valid_data_path, test_data_prefix=args.test_data_path)
    print_rank_0("> finished creating datasets ...")
    return train_ds, valid_ds, test_ds


def extra_args(parser):
    group = parser.add_argument_group(title="finetune")
    group.add_argument("--model_name", choices={"gpt", "llama", "falcon", "llama2", "codellama"}, default="gpt")
    group.add_argument("--model_type", choices={"encoder_or_decoder", "encoder_and_decoder"},
                       default="encoder_or_decoder")
    return parser


def main(args_list=None):
    args_defaults = {"tokenizer_type": "GPT2BPETokenizer"}
    initialize_megatron(extra_args, args_defaults, args_list=args_list)
    args = get_args()
    pretrain(args, train_valid_test_datasets_provider, model_provider, ModelType.encoder_or_decoder, forward_step)


if __name__ == "__main__":
    main()
