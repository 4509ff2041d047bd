"""Model zoo: the two workload families BASELINE.json:5 names (ResNet, BERT)."""
from .bert import BertConfig, BertForSequenceClassification, bert_base, bert_large, bert_tiny  # noqa: F401
from .resnet import ResNet, resnet18, resnet50  # noqa: F401


def build_model(name: str, **kw):
    table = {"resnet18": resnet18, "resnet50": resnet50, "bert-base": bert_base,
             "bert-large": bert_large, "bert-tiny": bert_tiny}
    if name not in table:
        raise KeyError(f"unknown model {name!r}; choose from {sorted(table)}")
    return table[name](**kw)
