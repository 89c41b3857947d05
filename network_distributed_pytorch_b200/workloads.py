"""Gradient shape sets of the BASELINE.json configurations (parameter order = model.parameters()).

Written out by hand so that benchmarks and tests do not need torchvision / transformers at run
time; tests/test_workloads.py checks them against the real model constructors where those
libraries are importable.
"""


def distilbert(vocab=30522, dim=768, hidden=3072, layers=6, max_pos=512, labels=2):
    """HF DistilBertForSequenceClassification(DistilBertConfig()): 40 matrices, 64 vectors."""
    s = [(vocab, dim), (max_pos, dim), (dim,), (dim,)]
    for _ in range(layers):
        for _ in range(4):  # q_lin, k_lin, v_lin, out_lin
            s += [(dim, dim), (dim,)]
        s += [(dim,), (dim,)]  # sa_layer_norm
        s += [(hidden, dim), (hidden,), (dim, hidden), (dim,)]  # ffn.lin1, ffn.lin2
        s += [(dim,), (dim,)]  # output_layer_norm
    s += [(dim, dim), (dim,), (labels, dim), (labels,)]  # pre_classifier, classifier
    return s


def bert(vocab=30522, dim=1024, hidden=4096, layers=24, max_pos=512, types=2, labels=2):
    """HF BertForSequenceClassification; defaults = BERT-large (149 matrices, 244 vectors)."""
    s = [(vocab, dim), (max_pos, dim), (types, dim), (dim,), (dim,)]
    for _ in range(layers):
        for _ in range(4):  # query, key, value, attention.output.dense
            s += [(dim, dim), (dim,)]
        s += [(dim,), (dim,)]  # attention.output.LayerNorm
        s += [(hidden, dim), (hidden,), (dim, hidden), (dim,)]  # intermediate, output
        s += [(dim,), (dim,)]  # output.LayerNorm
    s += [(dim, dim), (dim,), (labels, dim), (labels,)]  # pooler, classifier
    return s


def _bn(c):
    return [(c,), (c,)]


def resnet(depth=18, classes=1000):
    """torchvision resnet18 / resnet50 parameter shapes."""
    if depth == 18:
        block, counts, exp = "basic", [2, 2, 2, 2], 1
    elif depth == 34:
        block, counts, exp = "basic", [3, 4, 6, 3], 1
    elif depth == 50:
        block, counts, exp = "bottleneck", [3, 4, 6, 3], 4
    elif depth == 101:
        block, counts, exp = "bottleneck", [3, 4, 23, 3], 4
    elif depth == 152:
        block, counts, exp = "bottleneck", [3, 8, 36, 3], 4
    else:
        raise ValueError(depth)
    s = [(64, 3, 7, 7)] + _bn(64)
    inplanes = 64
    for stage, (planes, n) in enumerate(zip([64, 128, 256, 512], counts)):
        for b in range(n):
            stride = 2 if (stage > 0 and b == 0) else 1
            if block == "basic":
                s += [(planes, inplanes, 3, 3)] + _bn(planes) + [(planes, planes, 3, 3)] + _bn(planes)
            else:
                s += [(planes, inplanes, 1, 1)] + _bn(planes) + [(planes, planes, 3, 3)] + _bn(planes)
                s += [(planes * 4, planes, 1, 1)] + _bn(planes * 4)
            if b == 0 and (stride != 1 or inplanes != planes * exp):
                s += [(planes * exp, inplanes, 1, 1)] + _bn(planes * exp)
            inplanes = planes * exp
    s += [(classes, 512 * exp), (classes,)]
    return s


WORKLOADS = {
    "resnet18": lambda: resnet(18),
    "distilbert": distilbert,
    "resnet50": lambda: resnet(50),
    "bert-large": bert,
    "resnet152": lambda: resnet(152),
}


def stats(shapes, rank):
    """E_hi, E_P, E_Q, E_1 and the algorithmic byte count of SURVEY.md section 8(d)."""
    e_hi = e_p = e_q = e_1 = 0
    n_mat = n_vec = 0
    for s in shapes:
        numel = 1
        for d in s:
            numel *= d
        if len(s) <= 1:
            e_1 += numel
            n_vec += 1
        else:
            n = s[0]
            m = numel // n
            r = min(n, m, rank)
            e_hi += numel
            e_p += n * r
            e_q += m * r
            n_mat += 1
    return dict(n_mat=n_mat, n_vec=n_vec, e_hi=e_hi, e_p=e_p, e_q=e_q, e_1=e_1,
                grad_bytes=4 * (e_hi + e_1),
                algo_bytes=4 * (5 * e_hi + 4 * e_p + 3 * e_q + 4 * e_1),
                bits=32 * (e_p + e_q + e_1),
                k1_bytes=4 * (e_hi + e_q + e_p),
                k2_bytes=4 * (2 * e_p),
                k3_bytes=4 * (e_hi + e_p + e_q),
                k4_bytes=4 * (3 * e_hi + e_p + e_q),
                k6_bytes=4 * (4 * e_1))
