"""Data preparation for multi-stage dialogue prompting (tasks/msdp/preprocessing.py:12-581
surface): Wizard-of-Wikipedia / Wizard-of-Internet -> "topic \\t dialogue context \\t knowledge
\\t response" lines, prompt selection for the knowledge and response stages, and the input of the
response stage from generated knowledge.

    python tasks/msdp/preprocessing.py --func {process_wow_dataset,process_woi_dataset,
        get_knwl_gen_prompts,get_resp_gen_prompts,prepare_input} ...
"""
import argparse
import json
import random

import numpy as np


def get_args():
    p = argparse.ArgumentParser(description="Preprocessing")
    p.add_argument("--func", type=str, default=None)
    p.add_argument("--raw_file", type=str, default=None)
    p.add_argument("--processed_file", type=str, default=None)
    p.add_argument("--knwl_ref_file", type=str, default=None)
    p.add_argument("--resp_ref_file", type=str, default=None)
    p.add_argument("--knwl_gen_file", type=str, default=None)
    p.add_argument("--test_file", type=str, default=None)
    p.add_argument("--train_file", type=str, default=None)
    p.add_argument("--model_file", type=str, default=None, help="sentence encoder for prompt retrieval")
    p.add_argument("--data_type", type=str, default=None, choices=[None, "wow_seen", "wow_unseen", "woi"])
    p.add_argument("--seed", type=int, default=1234)
    return p.parse_args()


def _clean(s):
    return " ".join(s.replace("\n", " ").replace("\r", " ").replace("\t", " ").split())


def process_wow_dataset(raw_file, processed_file, knwl_ref_file, resp_ref_file):
    """WoW json -> one line per wizard turn; optional reference files for evaluation."""
    with open(raw_file) as f:
        dialogs = json.load(f)
    fk = open(knwl_ref_file, "w") if knwl_ref_file else None
    fr = open(resp_ref_file, "w") if resp_ref_file else None
    with open(processed_file, "w") as fout:
        for sample in dialogs:
            turns = []
            for i, item in enumerate(sample["dialog"]):
                text = _clean(item["text"])
                if "wizard" in item["speaker"].lower():
                    if i == 0:
                        turns.append(text)      # the wizard opens: nothing to predict yet
                        continue
                    checked = item.get("checked_sentence", {})
                    passage = item.get("checked_passage", {})
                    knowledge = _clean(list(checked.values())[0]) if len(checked) == 1 else "no_passages_used"
                    topic = _clean(list(passage.values())[0]) if len(passage) == 1 else sample["chosen_topic"]
                    if knowledge != "no_passages_used":
                        fout.write("\t".join([topic, " [SEP] ".join(turns), knowledge, text]) + "\n")
                        if fk:
                            fk.write(knowledge + "\n")
                        if fr:
                            fr.write(" ".join(text.split()) + "\n")
                turns.append(text)
    for f in (fk, fr):
        if f:
            f.close()


def process_woi_dataset(raw_file, processed_file, knwl_ref_file, resp_ref_file):
    """WoI jsonl -> same line format; the search query stands in for the topic."""
    fk = open(knwl_ref_file, "w") if knwl_ref_file else None
    fr = open(resp_ref_file, "w") if resp_ref_file else None
    with open(raw_file) as fin, open(processed_file, "w") as fout:
        for line in fin:
            for _, dialog in json.loads(line).items():
                turns, query = [], None
                for item in dialog["dialog_history"]:
                    action = item["action"]
                    if action == "Wizard => SearchAgent":
                        query = _clean(item["text"])
                    elif action == "Apprentice => Wizard":
                        turns.append(_clean(item["text"]))
                    elif action == "Wizard => Apprentice":
                        text = _clean(item["text"])
                        sel = item.get("context", {}).get("selected_contents", [])
                        contents = item.get("context", {}).get("contents", [])
                        picked = []
                        for flags, doc in zip(sel[1:], contents):
                            picked += [c for c, fl in zip(doc.get("content", []), flags) if fl]
                        knowledge = _clean(" ".join(picked)) if picked else "no_passages_used"
                        if turns and knowledge != "no_passages_used":
                            fout.write("\t".join([query or "no_topic", " [SEP] ".join(turns), knowledge, text]) + "\n")
                            if fk:
                                fk.write(knowledge + "\n")
                            if fr:
                                fr.write(" ".join(text.split()) + "\n")
                        turns.append(text)
    for f in (fk, fr):
        if f:
            f.close()


def get_database(test_datapath, train_datapath, data_type):
    """Training examples grouped by topic (dialogue examples for prompt selection)."""
    test_topics = set()
    with open(test_datapath) as f:
        for line in f:
            test_topics.add(line.split("\t")[0])
    by_topic, dialogs, examples = {}, [], []
    with open(train_datapath) as f:
        for line in f:
            parts = line.strip().split("\t")
            if len(parts) < 4 or parts[2] == "no_passages_used":
                continue
            topic, turns, knowledge = parts[0], parts[1].split(" [SEP] "), parts[2]
            last = turns[-1]
            if data_type != "wow_seen" and topic not in knowledge:
                continue
            ex = "( " + last + " ) " + topic + " => " + knowledge
            by_topic.setdefault(topic, []).append((last, ex))
            dialogs.append(topic + " " + last)
            examples.append(ex)
    return by_topic, dialogs, examples


def _encode(model, sentences):
    """Sentence embeddings; a bag-of-words hashing fallback keeps the tool usable offline."""
    if model is not None:
        return np.asarray(model.encode(sentences))
    dim = 512
    out = np.zeros((len(sentences), dim), dtype=np.float32)
    for i, s in enumerate(sentences):
        for w in s.lower().split():
            out[i, hash(w) % dim] += 1.0
        n = np.linalg.norm(out[i])
        if n:
            out[i] /= n
    return out


def select_prompts_based_on_similarity(query, dialog_list, example_list, topic, model, topk):
    q = _encode(model, [query])[0]
    d = _encode(model, dialog_list)
    order = np.argsort(-(d @ q))[:topk]
    return [example_list[i] for i in order][::-1]


def prompt_selection_for_knowledge_generation(test_datapath, train_datapath, model_path, output_prompt_path,
                                              data_type):
    model = None
    if model_path:
        try:
            from sentence_transformers import SentenceTransformer
            model = SentenceTransformer(model_path)
        except ImportError:
            print("sentence_transformers not available: using the hashing encoder")
    by_topic, dialogs, examples = get_database(test_datapath, train_datapath, data_type)
    with open(test_datapath) as fin, open(output_prompt_path, "w") as fout:
        for line in fin:
            parts = line.strip().split("\t")
            topic, last = parts[0], parts[1].split(" [SEP] ")[-1]
            key = topic + " " + last
            if topic in by_topic:
                pool = by_topic[topic]
                sel = select_prompts_based_on_similarity(key, [topic + " " + t for t, _ in pool],
                                                         [e for _, e in pool], topic, model, 10)
            else:
                sel = select_prompts_based_on_similarity(key, dialogs, examples, topic, model, 10)
            fout.write(json.dumps({key: sel}) + "\n")


def prompt_selection_for_response_generation(input_path, output_path, seed):
    random.seed(seed)
    prompts = []
    with open(input_path) as f:
        for line in f:
            parts = line.strip().split("\t")
            if len(parts) < 4 or parts[2] == "no_passages_used":
                continue
            topic, last, knowledge, resp = parts[0], parts[1].split(" [SEP] ")[-1], parts[2], parts[3]
            overlap = len(set(knowledge.lower().split()) & set(resp.lower().split()))
            if 0.6 * len(knowledge.split()) <= overlap and len(resp.split()) <= 45:
                prompts.append("Topic: " + topic + ". User says: " + last + " We know that: " + knowledge +
                               " System replies: " + resp)
    random.shuffle(prompts)
    with open(output_path, "w") as f:
        for p in prompts[:20]:
            f.write(p + "\n")


def prepare_input_for_response_generation(test_file, knwl_gen_file, processed_file):
    with open(knwl_gen_file) as f:
        knowledge = [ln.strip().replace("<|endoftext|>", "") for ln in f]
    with open(test_file) as fin, open(processed_file, "w") as fout:
        for i, line in enumerate(fin):
            parts = line.strip().split("\t")
            fout.write("\t".join([parts[0], parts[1], knowledge[i], parts[3] if len(parts) > 3 else ""]) + "\n")


if __name__ == "__main__":
    a = get_args()
    if a.func == "process_wow_dataset":
        process_wow_dataset(a.raw_file, a.processed_file, a.knwl_ref_file, a.resp_ref_file)
    elif a.func == "process_woi_dataset":
        process_woi_dataset(a.raw_file, a.processed_file, a.knwl_ref_file, a.resp_ref_file)
    elif a.func == "get_knwl_gen_prompts":
        prompt_selection_for_knowledge_generation(a.test_file, a.train_file, a.model_file, a.processed_file,
                                                  a.data_type)
    elif a.func == "get_resp_gen_prompts":
        prompt_selection_for_response_generation(a.train_file, a.processed_file, a.seed)
    elif a.func == "prepare_input":
        prepare_input_for_response_generation(a.test_file, a.knwl_gen_file, a.processed_file)
