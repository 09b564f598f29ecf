"""RACE finetuning on the BERT multiple-choice head (tasks/race/finetune.py:14-58 surface)."""
from epfl_megatron_b200 import get_args, get_tokenizer, print_rank_0
from epfl_megatron_b200.models.multiple_choice import MultipleChoice
from eval_utils import accuracy_func_provider
from finetune_utils import finetune
from race.data import RaceDataset


def train_valid_datasets_provider():
    args, tok = get_args(), get_tokenizer()
    return (RaceDataset("training", args.train_data, tok, args.seq_length),
            RaceDataset("validation", args.valid_data, tok, args.seq_length))


def model_provider(pre_process=True, post_process=True):
    print_rank_0("building multichoice model for RACE ...")
    return MultipleChoice(num_tokentypes=2, pre_process=pre_process, post_process=post_process)


def metrics_func_provider():
    args, tok = get_args(), get_tokenizer()

    def single_dataset_provider(datapath):
        return RaceDataset(datapath.split("RACE")[-1].strip("/").replace("/", "-"), [datapath], tok,
                           args.seq_length)
    return accuracy_func_provider(single_dataset_provider)


def main():
    finetune(train_valid_datasets_provider, model_provider, end_of_epoch_callback_provider=metrics_func_provider)
