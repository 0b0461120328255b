import torch, sys
sys.path.insert(0, '.')
from commefficient_b200 import models
from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
from commefficient_b200.gpt2_train import compute_loss_train, compute_loss_val
from commefficient_b200.utils import make_args
from commefficient_b200.utils import flat
orig = flat.FlatParams.range_final
calls = []
def spy(self, lo, hi):
    calls.append((lo, hi, self.on_range_final is not None))
    return orig(self, lo, hi)
flat.FlatParams.range_final = spy
W, lb, seq = 8, 8, 128
args = make_args(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9, num_workers=W,
                 num_clients=17568, local_batch_size=lb, dataset_name="PERSONA", device="cuda", num_results_train=1,
                 weight_decay=0.0, precision="bf16", no_cuda_graphs=True)
torch.manual_seed(0)
model = models.GPT2DoubleHeadsModel(); model.resize_token_embeddings(50257 + 5)
fed = FedModel(model, compute_loss_train, args, compute_loss_val)
opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args); opt.param_groups[0]["lr"] = 0.01
g = torch.Generator().manual_seed(1); n = W * lb
ids = torch.randint(0, 50257, (n, 2, seq), generator=g); labels = ids.clone(); labels[:, 0] = -1; labels[:, 1, :seq // 2] = -1
cid = torch.arange(W).repeat_interleave(lb).long()
batch = (cid, ids, torch.full((n, 2), seq - 1), labels, torch.randint(0, 2, (n,), generator=g), torch.randint(0, 50257, (n, 2, seq), generator=g))
fed.train(True)
for i in range(2):
    calls.clear()
    out = fed(batch); opt.step()
    torch.cuda.synchronize()
    print("round", i, "range_final calls", len(calls), calls[:3], "d", fed.fp.numel)
