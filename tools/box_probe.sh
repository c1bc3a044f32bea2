#!/bin/bash
# What the GPU box looks like: GPUs, PCIe topology, host cores and memory.  Output goes to gpurun_out/box_probe.txt.
out=gpurun_out/box_probe.txt
{
  echo "== nvidia-smi -L"; nvidia-smi -L
  echo "== nvidia-smi topo -m"; nvidia-smi topo -m
  echo "== nproc / memory"; nproc; free -g | head -2
  echo "== lscpu"; lscpu | grep -E 'Model name|Socket|NUMA|Thread|Core|^CPU\(s\)'
  echo "== pcie link"; nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current,pcie.link.width.max --format=csv
  echo "== go"; (go version 2>&1 || true)
} > $out 2>&1
cat $out
