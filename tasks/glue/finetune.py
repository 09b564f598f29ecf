"""GLUE finetuning: MNLI (3 classes) and QQP (2 classes) on the BERT classifier
(tasks/glue/finetune.py:14-93 surface)."""
from epfl_megatron_b200 import get_args, get_tokenizer, print_rank_0
from epfl_megatron_b200.models.classification import Classification
from eval_utils import accuracy_func_provider
from finetune_utils import finetune


def glue_classification(num_classes, Dataset, name_from_datapath_func):
    def train_valid_datasets_provider():
        args, tok = get_args(), get_tokenizer()
        return (Dataset("training", args.train_data, tok, args.seq_length),
                Dataset("validation", args.valid_data, tok, args.seq_length))

    def model_provider(pre_process=True, post_process=True):
        print_rank_0(f"building classification model for {get_args().task} ...")
        return Classification(num_classes=num_classes, num_tokentypes=2, pre_process=pre_process,
                              post_process=post_process)

    def metrics_func_provider():
        args, tok = get_args(), get_tokenizer()

        def single_dataset_provider(datapath):
            return Dataset(name_from_datapath_func(datapath), [datapath], tok, args.seq_length)
        return accuracy_func_provider(single_dataset_provider)

    finetune(train_valid_datasets_provider, model_provider, end_of_epoch_callback_provider=metrics_func_provider)


def main():
    args = get_args()
    if args.task == "MNLI":
        from glue.mnli import MNLIDataset as Dataset

        def name_from_datapath(p):
            return p.split("MNLI")[-1].strip(".tsv").strip("/").replace("_", "-")
        glue_classification(3, Dataset, name_from_datapath)
    elif args.task == "QQP":
        from glue.qqp import QQPDataset as Dataset

        def name_from_datapath(p):
            return p.split("QQP")[-1].strip(".tsv").strip("/").replace("_", "-")
        glue_classification(2, Dataset, name_from_datapath)
    else:
        raise NotImplementedError(f"GLUE task {args.task} is not implemented.")
