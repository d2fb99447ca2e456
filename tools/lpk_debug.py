import os, sys, traceback, collections
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "torchrec-oldfork_b200")]
import bench
from torchrec.sparse import jagged_tensor as jt
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
sys.argv = [sys.argv[0]]; args = bench.parse_args()
dmp, opt, rows = bench.build_gpu_model(args, rank, 2, dev, dist.group.WORLD)
host = bench.make_batches(rows, args.batch, rank, 2)
for hb in host: hb.sparse_features.length_per_key()
batches = [b.to(dev) for b in host]
import itertools
from torchrec.distributed.train_pipeline import TrainPipelineSparseDist
pipe = TrainPipelineSparseDist(dmp, opt, dev); it = itertools.cycle(batches)
for i in range(3): pipe.progress(it)
orig = jt.KeyedJaggedTensor.length_per_key
sites = collections.Counter()
def patched(self):
    if self._length_per_key is None and len(self._keys):
        st = traceback.extract_stack(limit=6)
        sites[" <- ".join(f"{os.path.basename(f.filename)}:{f.lineno}" for f in st[:-1][::-1][:4])] += 1
    return orig(self)
jt.KeyedJaggedTensor.length_per_key = patched
for i in range(4): pipe.progress(it)
torch.cuda.synchronize()
if rank == 0:
    for k, v in sites.most_common(): print(v, k)
dist.destroy_process_group()
