"""Drive the UNMODIFIED upstream training scripts (``ddp_init.setup(); run_task(); cleanup()``) for a few batches with
either this package's drop-in ``reducer`` / ``tensor_buffer`` modules or the reference's own, and dump what came out.

Run as a script in a fresh process (module state, the default process group and ``sys.path`` are all global):

    python tests/upstream_loop.py <cifar|imdb> <dropin|reference> <out.npz> [cpu|cuda]

Nothing of the upstream file is edited.  What is patched around it is what SURVEY.md section 7 step 5 lists, because
none of it is reachable offline: the rendezvous (the CIFAR script hard-codes a ``tcp://`` literal inside the call,
ddp_init.py:94 -> a ``file://`` rendezvous on gloo), the dataset (``datasets.CIFAR10(download=True)`` / the IMDb files
under an absolute home path -> seeded synthetic samples), the model (``resnet152(pretrained=True)`` /
``from_pretrained('distilbert-base-uncased')`` -> a tiny seeded model of the same family) and ``transformers.AdamW``
(imported, never used, removed from transformers 5).  On CPU the engine of the drop-in is the oracle-backed test
double of tests/fake_engine.py (the product engine needs a GPU); on CUDA it is the real library.
"""
import importlib.util
import os
import sys
import tempfile
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
SCRIPT_DIRS = {"cifar": "ddp_powersgd_guide_cifar10", "imdb": "ddp_powersgd_distillBERT_IMDb"}
DROPIN = os.path.join(ROOT, "network_distributed_pytorch_b200", "dropin")


def _tiny_cnn():
    torch.manual_seed(99)
    return torch.nn.Sequential(
        torch.nn.Conv2d(3, 8, 3, stride=2, padding=1), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
        torch.nn.Conv2d(8, 16, 3, stride=2, padding=1), torch.nn.BatchNorm2d(16), torch.nn.ReLU(),
        torch.nn.AdaptiveAvgPool2d(2), torch.nn.Flatten(), torch.nn.Linear(64, 10))


class _FakeCifar(torch.utils.data.Dataset):
    """Stands in for torchvision.datasets.CIFAR10: same constructor keywords, seeded random images."""

    def __init__(self, root, train=True, download=False, transform=None):
        g = torch.Generator().manual_seed(7)
        self.x = torch.rand(1024, 3, 32, 32, generator=g) * 2 - 1
        self.y = torch.randint(0, 10, (1024,), generator=g)

    def __len__(self):
        return len(self.y)

    def __getitem__(self, i):
        return self.x[i], int(self.y[i])


def main(which, reducer_kind, out_path, device="cpu"):
    warnings.filterwarnings("ignore")
    torch.set_num_threads(2)
    script_dir = os.path.join(REF, SCRIPT_DIRS[which])
    sys.path.insert(0, ROOT)
    # flat module names resolve through sys.path (ddp_init.py:20, reducer.py:4): drop-in first, or the reference's own
    sys.path.insert(0, script_dir)  # partition_helper (and, for "reference", reducer / tensor_buffer)
    if reducer_kind == "dropin":
        sys.path.insert(0, DROPIN)
        if device == "cpu":  # test-side engine swap: the product engine refuses CPU devices
            import network_distributed_pytorch_b200.engine as engine_mod
            from tests.fake_engine import OracleEngine

            engine_mod.CudaEngine = OracleEngine

    import transformers

    if not hasattr(transformers, "AdamW"):
        transformers.AdamW = torch.optim.AdamW  # ddp_init.py:16 imports the symbol and never uses it

    spec = importlib.util.spec_from_file_location("_upstream_ddp_init_" + which, os.path.join(script_dir, "ddp_init.py"))
    ddp_init = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ddp_init)

    import reducer as reducer_mod
    import tensor_buffer as tb_mod

    expect = DROPIN if reducer_kind == "dropin" else script_dir
    assert os.path.dirname(os.path.abspath(reducer_mod.__file__)) == expect, reducer_mod.__file__
    assert os.path.dirname(os.path.abspath(tb_mod.__file__)) == expect, tb_mod.__file__
    assert ddp_init.PowerSGDReducer is reducer_mod.PowerSGDReducer

    tmp = tempfile.mkdtemp()
    rdzv = "file://" + os.path.join(tmp, "rdzv")
    cfg = ddp_init.config
    cfg.update(rank=0, cuda_rank=0, n_workers=1, distributed_backend="gloo", training_epochs=1, output_dir=tmp,
               reducer_rank=2)
    if "init_method" in cfg:
        cfg["init_method"] = rdzv
    real_init = torch.distributed.init_process_group

    def init_with_local_rendezvous(*a, **kw):
        kw["init_method"] = rdzv  # the CIFAR script passes a tcp:// literal
        return real_init(*a, **kw)

    ddp_init.dist.init_process_group = init_with_local_rendezvous
    if device == "cpu":
        torch.cuda.is_available = lambda: False  # the scripts pick their device from this (ddp_init.py:71,107)

    made = {}
    if which == "cifar":
        ddp_init.datasets.CIFAR10 = _FakeCifar

        def tiny_resnet(pretrained=True):
            made["model"] = _tiny_cnn()
            return made["model"]

        ddp_init.models.resnet152 = tiny_resnet
    else:
        def synthetic_imdb():
            g = torch.Generator().manual_seed(7)
            n, seq = 64, 32
            enc = {"input_ids": torch.randint(0, 1000, (n, seq), generator=g).tolist(),
                   "attention_mask": torch.ones(n, seq, dtype=torch.long).tolist()}
            labels = torch.randint(0, 2, (n,), generator=g).tolist()
            return ddp_init.partition_dataset(ddp_init.IMDbDataset(enc, labels)), None, None

        ddp_init.prepare_IMDb = synthetic_imdb

        class TinyDistilBert:
            @staticmethod
            def from_pretrained(name):
                cfgm = transformers.DistilBertConfig(vocab_size=1000, dim=32, hidden_dim=64, n_layers=1, n_heads=2,
                                                     max_position_embeddings=32, dropout=0.0, attention_dropout=0.0,
                                                     seq_classif_dropout=0.0)
                torch.manual_seed(99)
                made["model"] = transformers.DistilBertForSequenceClassification(cfgm)
                return made["model"]

        ddp_init.DistilBertForSequenceClassification = TinyDistilBert

    # record what the loop hands to and gets from the reducer
    calls = {"n": 0, "bits": []}
    real_reduce = reducer_mod.PowerSGDReducer.reduce

    def counting_reduce(self, grad_in, grad_out, memory_out):
        bits = real_reduce(self, grad_in, grad_out, memory_out)
        calls["n"] += 1
        calls["bits"].append(int(bits))
        return bits

    reducer_mod.PowerSGDReducer.reduce = counting_reduce

    ddp_init.setup()
    ddp_init.run_task()
    ddp_init.cleanup()

    params = torch.cat([p.detach().reshape(-1).cpu() for p in made["model"].parameters()]).numpy()
    np.savez(out_path, params=params, calls=np.int64(calls["n"]), bits=np.array(calls["bits"], np.int64),
             reducer_file=np.array(os.path.abspath(reducer_mod.__file__)))


if __name__ == "__main__":
    main(*sys.argv[1:])
