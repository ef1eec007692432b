"""Prints how many CTA pairs / split-K clusters of 2 the device holds at once (cudaOccupancyMaxActiveClusters)."""
import sys
sys.path.insert(0, ".")
import torch
from int_act_b200 import binding as pb
torch.zeros(1, device="cuda")
print("max pairs", pb.lib().pi0_debug_max_pairs(), "max clusters(2) splitk", pb.lib().pi0_debug_max_clusters(2))
