import sys, torch
sys.path.insert(0, "/root/repo")
import bench, mast3r_slam_backends as be
dev = torch.device("cuda:0")
print(bench.run_other_configs(dev, be, 10))
