"""The HBM-bound rows alone (bench.hbm_rows): a15 / a16 / a17 of the covariance estimator at C2 size and the RFF of C3."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
lib = importlib.import_module(bench.PKG + "._lib").load()
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
print(json.dumps(bench.hbm_rows(torch, dev, lib), indent=1))
