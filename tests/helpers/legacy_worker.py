"""Subprocess body of tests/test_legacy_models.py: builds a toy WordPiece corpus, then trains a
tiny BERT / T5 / ICT biencoder for a few steps on CPU through the real entry-point functions."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

WORDS = ("the of and to in a is that for it as was with be by on not he this are or his from at which but have "
         "an had they you were their one all we can her has there been if more when will would who so no").split()


def make_corpus(d):
    import numpy as np
    from epfl_megatron_b200.data import indexed_dataset
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + WORDS + ["##s", "##ed", "##ing", ".", ","]
    vf = os.path.join(d, "vocab.txt")
    with open(vf, "w") as f:
        f.write("\n".join(vocab) + "\n")
    rng = np.random.RandomState(0)
    prefix = os.path.join(d, "corpus_text_sentence")
    b = indexed_dataset.make_builder(prefix + ".bin", impl="mmap", vocab_size=len(vocab))
    tb = indexed_dataset.make_builder(os.path.join(d, "titles") + ".bin", impl="mmap", vocab_size=len(vocab))
    for _ in range(40):                       # documents
        for _ in range(rng.randint(3, 7)):    # sentences
            n = rng.randint(5, 14)
            b.add_item(np.asarray(rng.randint(5, 5 + len(WORDS), size=n), dtype=np.int64))
        b.end_document()
        tb.add_item(np.asarray(rng.randint(5, 5 + len(WORDS), size=3), dtype=np.int64))
        tb.end_document()
    b.finalize(prefix + ".idx")
    tb.finalize(os.path.join(d, "titles") + ".idx")
    return vf, prefix, os.path.join(d, "titles")


def main(mode, workdir):
    import torch
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    vf, prefix, titles = make_corpus(workdir)
    from epfl_megatron_b200 import get_args
    from epfl_megatron_b200.initialize import initialize_megatron
    from epfl_megatron_b200.models.enums import ModelType
    from epfl_megatron_b200.training import (build_train_valid_test_data_iterators, setup_model_and_optimizer,
                                             train_step)
    common = (f"--num_layers 2 --hidden_size 64 --num_attention_heads 4 --seq_length 48 "
              f"--max_position_embeddings 64 --micro_batch_size 4 --global_batch_size 4 --train_iters 6 "
              f"--lr 2e-3 --min_lr 1e-4 --lr_decay_style constant --tokenizer_type BertWordPieceLowerCase "
              f"--vocab_file {vf} --data_path {prefix} --split 90,5,5 --distributed_backend gloo "
              f"--use_cpu_initialization --hidden_dropout 0.0 --attention_dropout 0.0 --no_bias_gelu_fusion "
              f"--no_masked_softmax_fusion --num_workers 0 --eval_iters 1 --eval_interval 100 "
              f"--log_interval 100 --make_vocab_size_divisible_by 8 --data_impl mmap --seed 7").split()
    if mode == "bert":
        import pretrain_bert as ep
        argv, mt, provider = common, ModelType.encoder_or_decoder, ep.model_provider
    elif mode == "t5":
        import pretrain_t5 as ep
        argv = common + "--decoder_seq_length 32 --vocab_extra_ids 100 --kv_channels 16 --ffn_hidden_size 128".split()
        mt, provider = ModelType.encoder_and_decoder, ep.model_provider
    else:
        import pretrain_ict as ep
        argv = common + (f"--titles_data_path {titles} --query_in_block_prob 0.2 --ict_head_size 32 "
                         "--biencoder_projection_dim 32 --retriever_report_topk_accuracies 1 2 "
                         "--retriever_score_scaling").split()
        mt, provider = ModelType.encoder_or_decoder, ep.pretrain_ict_model_provider
    initialize_megatron(None, {}, args_list=argv)
    args = get_args()
    model, optimizer, sched = setup_model_and_optimizer(provider, mt)
    train_it, _, _ = build_train_valid_test_data_iterators(ep.train_valid_test_datasets_provider)
    for m in model:
        m.train()
    losses = []
    for _ in range(6):
        ld, skipped, gn, _ = train_step(ep.forward_step, train_it, model, optimizer, sched, args)
        key = "lm loss" if "lm loss" in ld else "loss"
        losses.append(float(ld[key]))
    # checkpoint-format round trip of the model's own state dict
    sd = model[0].state_dict_for_save_checkpoint() if hasattr(model[0], "state_dict_for_save_checkpoint") \
        else model[0].module.state_dict_for_save_checkpoint()
    print("RESULT " + json.dumps({"losses": losses, "keys": sorted(sd.keys())}))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
