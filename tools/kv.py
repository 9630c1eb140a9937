"""Experiment helper: the stride-2 / downsample step times of ResNet18 from a bench_detail.json."""
import json,sys
d=json.load(open(sys.argv[1]))['models']['resnet18']
k=d['kernel_ms']
print(sys.argv[1], 'step', round(d['ms_per_step'],4), {n:k[n] for n in k if '.0.conv1' in n or 'downsample' in n})
