"""Host-side restatement of the upstream PowerSGD training loop (``run_task`` in
ddp_powersgd_distillBERT_IMDb/ddp_init.py:140-231, same structure in the CIFAR script) on synthetic data, used to
report training samples/sec with this reducer dropped in.  Two forms:

* ``loop="upstream"``  -- line for line what the upstream loop does around the reducer: per-parameter
  ``send = g + e`` (ddp_init.py:199-201), ``reducer.reduce(send, grads, memories)`` (:205), per-parameter momentum
  and parameter update (:210-222).  Only the reducer is ours.
* ``loop="fused"``     -- the same arithmetic with the per-parameter loops folded into kernels:
  ``reducer.reduce_fused`` and ``FusedMomentumSGD.step``.

Datasets, pretrained weights and rendezvous addresses of the upstream scripts are not reachable offline, so the
model is randomly initialised from its config and the batches are random tokens of the upstream shapes
(batch 16 per GPU, 512 tokens, ddp_init.py:75,89-90).
"""
import time

import torch


def build_model(name, device, dropout=None, num_labels=2):
    if name in ("distilbert", "distilbert-tiny", "bert-large"):
        import transformers

        if name == "bert-large":
            cfg = transformers.BertConfig(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
                                          intermediate_size=4096)
            if dropout is not None:
                cfg.hidden_dropout_prob = cfg.attention_probs_dropout_prob = dropout
            model = transformers.BertForSequenceClassification(cfg)
        else:
            cfg = transformers.DistilBertConfig()
            if name == "distilbert-tiny":
                cfg = transformers.DistilBertConfig(vocab_size=1000, dim=64, hidden_dim=128, n_layers=2, n_heads=4,
                                                    max_position_embeddings=64)
            if dropout is not None:
                cfg.dropout = cfg.attention_dropout = cfg.seq_classif_dropout = dropout
            cfg.num_labels = num_labels
            model = transformers.DistilBertForSequenceClassification(cfg)
        vocab, kind = cfg.vocab_size, "tokens"
    elif name in ("resnet18", "resnet50"):
        import torchvision

        model = getattr(torchvision.models, name)()
        vocab, kind = None, "images"
    else:
        raise ValueError(name)
    return model.to(device).train(), vocab, kind


def make_batch(kind, vocab, batch, seq_len, device, gen, num_labels=2):
    if kind == "tokens":
        return dict(input_ids=torch.randint(0, vocab, (batch, seq_len), device=device, generator=gen),
                    attention_mask=torch.ones(batch, seq_len, dtype=torch.long, device=device),
                    labels=torch.randint(0, num_labels, (batch,), device=device, generator=gen))
    size = 32 if seq_len <= 64 else 224
    return dict(x=torch.rand(batch, 3, size, size, device=device, generator=gen) * 2 - 1,
                y=torch.randint(0, 10, (batch,), device=device, generator=gen))


def train_steps(model, kind, batches, reducer, steps, lr=0.001, momentum=0.9, loop="upstream", autocast=True,
                state=None):
    """Run `steps` iterations of Algorithm 2 (error-feedback SGD with momentum).  Returns the state dict
    (memories, send buffers, momenta, step counter) so that calls can be chained."""
    params = [p for p in model.parameters()]
    if state is None:
        state = dict(memories=[torch.zeros_like(p) for p in params],          # ddp_init.py:174
                     send=[torch.zeros_like(p) for p in params],              # :176
                     momenta=[torch.empty_like(p) for p in params],           # :178
                     i=0, fused_opt=None, losses=[])
    memories, send, momenta = state["memories"], state["send"], state["momenta"]
    criterion = torch.nn.CrossEntropyLoss()
    if loop == "fused" and state["fused_opt"] is None:
        from .train_step import FusedMomentumSGD

        state["fused_opt"] = FusedMomentumSGD(params, lr, momentum)
    for k in range(steps):
        batch = batches[state["i"] % len(batches)]
        for p in params:                      # optimizer.zero_grad() (set_to_none), ddp_init.py:185
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            if kind == "tokens":
                loss = model(batch["input_ids"], attention_mask=batch["attention_mask"], labels=batch["labels"])[0]
            else:
                loss = criterion(model(batch["x"]), batch["y"])
        state["losses"].append(loss.item())   # epoch_loss += loss.item(), :191
        loss.backward()
        grads = [p.grad for p in params]      # :195
        if loop == "upstream":
            for g, e, s in zip(grads, memories, send):            # :199-201
                s.data[:] = g + e
            reducer.reduce(send, grads, memories)                 # :205
            for g, m in zip(grads, momenta):                      # :210-216
                if state["i"] == 0:
                    m.data = g.clone().detach()
                else:
                    m.mul_(momentum).add_(g)
                g.data[:] += m
            for p, g in zip(params, grads):                       # :221-222
                p.data.add_(g, alpha=-lr)
        else:
            reducer.reduce_fused(grads, memories, send)
            state["fused_opt"].step([p.data for p in params], grads, momenta)
        state["i"] += 1
    return state


def measure_samples_per_sec(name, reducer_factory, device, batch=16, seq_len=512, steps=10, warmup=3, loop="upstream",
                            world=1, barrier=None):
    """samples/sec of the whole job (world * batch per step), device-timed, max over ranks left to the caller."""
    model, vocab, kind = build_model(name, device)
    gen = torch.Generator(device=device).manual_seed(4321)
    batches = [make_batch(kind, vocab, batch, seq_len, device, gen) for _ in range(4)]
    reducer = reducer_factory()
    state = train_steps(model, kind, batches, reducer, warmup, loop=loop)
    if barrier:
        barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    train_steps(model, kind, batches, reducer, steps, loop=loop, state=state)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ms = e0.elapsed_time(e1)
    return dict(ms_per_step=ms / steps, wall_ms_per_step=wall * 1e3 / steps, batch_per_gpu=batch, seq_len=seq_len,
                steps=steps, warmup=warmup, loop=loop, model=name, final_loss=state["losses"][-1])
