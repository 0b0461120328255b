"""The benchmark configurations of BASELINE.json as plain data, shared by both arms (``bench.py`` for this
framework, ``baseline/run_ref.py`` for the unmodified reference).  No imports of either code base.

Every entry: ``family`` ("cv" | "gpt2"), the model and its constructor arguments, the synthetic input shape, the
federated flags, and the round geometry: ``clients_per_gpu`` x ``per_client`` samples per client (weak scaling) --
``--scaling strong`` keeps ``strong_clients`` clients per round whatever the GPU count.
"""

CONFIGS = {
    # BASELINE.json config 2 (headline): ResNet-9, CIFAR-10-shaped, FetchSGD sketch 5 x 500 000, k = 50 000
    "resnet9_sketch": dict(
        family="cv", model="ResNet9", model_kw=dict(num_classes=10), shape=(3, 32, 32), classes=10, dataset="CIFAR10",
        flags=dict(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9),
        clients_per_gpu=100, per_client=5, num_clients=10000, strong_clients=96),
    # config 3: ResNet-18, CIFAR-100-shaped, true_topk / local_topk with error accumulation
    "resnet18_true_topk": dict(
        family="cv", model="ResNet18", model_kw=dict(num_classes=100), shape=(3, 32, 32), classes=100, dataset="CIFAR100",
        flags=dict(mode="true_topk", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9, k=50000),
        clients_per_gpu=16, per_client=8, num_clients=10000, strong_clients=96),
    "resnet18_local_topk": dict(
        family="cv", model="ResNet18", model_kw=dict(num_classes=100), shape=(3, 32, 32), classes=100, dataset="CIFAR100",
        flags=dict(mode="local_topk", error_type="local", local_momentum=0.9, virtual_momentum=0.0, k=50000),
        # (the reference keeps num_clients x d floats of velocity AND of error in host shared memory:
        #  1024 clients x 5.25 M x 4 B x 2 = 43 GB is what fits this box; ours shards that state over the ranks)
        clients_per_gpu=16, per_client=8, num_clients=1024, strong_clients=96),
    # config 4: GPT-2 small double-heads, PersonaChat-shaped, sketch with virtual momentum
    "gpt2_sketch": dict(
        family="gpt2", model="GPT2DoubleHeadsModel", model_kw={}, seq=128, candidates=2, dataset="PERSONA",
        flags=dict(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9, weight_decay=0.0),
        clients_per_gpu=8, per_client=8, num_clients=17568, strong_clients=16),
    # config 5: FEMNIST-shaped ResNet-9 (1 x 28 x 28, 62 classes, 3500 clients): fedavg vs uncompressed vs sketch
    "femnist_sketch": dict(
        family="cv", model="ResNet9", model_kw=dict(num_classes=62, initial_channels=1), shape=(1, 28, 28), classes=62,
        dataset="EMNIST", flags=dict(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9),
        clients_per_gpu=32, per_client=20, num_clients=3500, strong_clients=96),
    "femnist_uncompressed": dict(
        family="cv", model="ResNet9", model_kw=dict(num_classes=62, initial_channels=1), shape=(1, 28, 28), classes=62,
        dataset="EMNIST", flags=dict(mode="uncompressed", error_type="none", local_momentum=0.0, virtual_momentum=0.9),
        clients_per_gpu=32, per_client=20, num_clients=3500, strong_clients=96),
    "femnist_fedavg": dict(
        family="cv", model="ResNet9", model_kw=dict(num_classes=62, initial_channels=1), shape=(1, 28, 28), classes=62,
        dataset="EMNIST", flags=dict(mode="fedavg", error_type="none", local_momentum=0.0, virtual_momentum=0.0,
                                     num_fedavg_epochs=2, fedavg_batch_size=10),
        clients_per_gpu=32, per_client=20, num_clients=3500, strong_clients=96),
}
# The same FEMNIST task on images zero-padded to 32 x 32, the size the reference's ResNet-9 can process (on 28 x 28
# its final MaxPool2d(4) meets a 3 x 3 map and raises): the configurations both arms can run.
for _name in ("sketch", "uncompressed", "fedavg"):
    _c = dict(CONFIGS["femnist_" + _name])
    _c["shape"] = (1, 32, 32)
    CONFIGS["femnist32_" + _name] = _c


def clients_per_round(cfg, gpus, scaling):
    """Weak: clients_per_gpu x gpus.  Strong: a fixed round (a multiple of every worker count the reference uses
    on 1 / 2 / 4 / 8 GPUs -- 1, 1, 3, 7 worker processes -- would be 21; 96 clients divide by 1, 2, 3, 4, 8 and the
    reference's remainder clients are reported as processed only when they were)."""
    return cfg["clients_per_gpu"] * gpus if scaling == "weak" else cfg["strong_clients"]


def cv_round(cfg, clients, seed, torch, noise=0.5):
    """One round's (client_ids, images, labels) of class-prototype + noise images (both arms draw identical data)."""
    g = torch.Generator().manual_seed(seed)
    shape, classes, per = tuple(cfg["shape"]), cfg["classes"], cfg["per_client"]
    protos = torch.randn((classes,) + shape, generator=torch.Generator().manual_seed(1234))
    n = clients * per
    labels = torch.randint(0, classes, (n,), generator=g)
    images = protos[labels] + noise * torch.randn((n,) + shape, generator=g)
    first = (seed * clients) % max(1, cfg["num_clients"] - clients)
    ids = (torch.arange(clients).repeat_interleave(per) + first).long()
    return ids, images, labels


def gpt2_round(cfg, clients, seed, torch, vocab=50257):
    """PersonaChat-shaped synthetic batch in the reference's tuple order
    (client_id, input_ids, mc_token_ids, lm_labels, mc_labels, token_type_ids); only the reply half of the gold
    candidate carries LM labels."""
    g = torch.Generator().manual_seed(1000 + seed)
    per, seq, cand = cfg["per_client"], cfg["seq"], cfg["candidates"]
    n = clients * per
    ids = torch.randint(0, vocab, (n, cand, seq), generator=g)
    labels = ids.clone()
    labels[:, :-1] = -1
    labels[:, -1, : seq // 2] = -1
    first = (seed * clients) % max(1, cfg["num_clients"] - clients)
    cid = (torch.arange(clients).repeat_interleave(per) + first).long()
    return (cid, ids, torch.full((n, cand), seq - 1), labels, torch.randint(0, cand, (n,), generator=g),
            torch.randint(0, vocab, (n, cand, seq), generator=g))
