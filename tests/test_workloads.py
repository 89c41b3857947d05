"""The hand-written gradient shape sets against the real model constructors (when importable)."""
import pytest

from network_distributed_pytorch_b200 import workloads


def test_survey_numbers():
    # SURVEY.md section 8(a-5), (a-8), (a-14) and BASELINE.md section 3
    st = workloads.stats(workloads.distilbert(), 4)
    assert (st["n_mat"], st["n_vec"], st["e_p"], st["e_q"], st["e_1"]) == (40, 64, 293100, 176640, 62210)
    assert st["bits"] == 17022400
    st = workloads.stats(workloads.resnet(18), 2)
    assert (st["n_mat"], st["n_vec"], st["e_p"], st["e_q"], st["e_1"], st["bits"]) == (21, 41, 11600, 63270, 10600, 2735040)
    st = workloads.stats(workloads.bert(), 8)
    assert (st["n_mat"], st["n_vec"], st["e_p"], st["e_q"], st["e_1"]) == (149, 244, 2025944, 1798144, 322562)
    st = workloads.stats(workloads.resnet(50), 4)
    assert (st["n_mat"], st["n_vec"], st["e_p"], st["e_q"], st["e_1"]) == (54, 107, 110240, 219724, 54120)


def test_against_torchvision():
    models = pytest.importorskip("torchvision.models")
    for depth in (18, 50):
        m = getattr(models, f"resnet{depth}")()
        assert [tuple(p.shape) for p in m.parameters()] == workloads.resnet(depth)


def test_against_transformers():
    tr = pytest.importorskip("transformers")
    m = tr.DistilBertForSequenceClassification(tr.DistilBertConfig())
    assert [tuple(p.shape) for p in m.parameters()] == workloads.distilbert()
    cfg = tr.BertConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128)
    m = tr.BertForSequenceClassification(cfg)
    assert [tuple(p.shape) for p in m.parameters()] == workloads.bert(dim=64, hidden=128, layers=2)
