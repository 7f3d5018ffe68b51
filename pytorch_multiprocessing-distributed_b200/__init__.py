"""B200-native multiprocess data-parallel training framework.

Import it as ``pmdist`` (see ``pmdist/__init__.py``): this directory carries the
prescribed project name, which contains a hyphen and so cannot be imported directly.

Layout
------
models/    CIFAR ResNet-18/34/50/101/152 with the reference's state_dict keys
ops/       ctypes bindings to the hand-written sm_100a kernels in ``csrc/``
parallel/  launcher, rendezvous, symmetric peer memory, bucket planner, step engine
data/      DistributedSampler-exact sharding, synthetic + CIFAR-10 sources, GPU augmentation
utils/     AverageMeter, Logger, accuracy, LR table, checkpoint, plot_curves
cli.py     flag system (superset of the reference's argparse flags)
train.py   per-rank trainer: epoch loop, rank-0 logs, checkpoint, plot
"""
