"""Shared pieces of the four example scripts.

The examples mirror the reference's four entry points (same flags, same prints, same hyper-parameters):
  convnet_spawn.py   <- mpspawn_dist.py   (ConvNet / MNIST, mp.spawn, env:// with explicit rank)
  convnet_launch.py  <- launch_dist.py    (ConvNet / MNIST, launcher env contract)
  resnet_spawn.py    <- example_mp.py     (ResNet-18 / CIFAR-10, mp.spawn, tcp:// init)
  resnet_launch.py   <- example_launch.py (ResNet-18 / CIFAR-10, launcher)
Only the imports differ: everything comes from this package instead of torch.distributed / torch.nn /
torchvision.  The training boxes have no network, so datasets are synthetic by default; pass --data-root to
read MNIST (IDX files) / CIFAR-10 (python pickles) from disk with this package's own readers and transforms.
"""
from __future__ import annotations

import os
import sys
from datetime import datetime

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
import pytorch_distributed_learning_b200.nn as nn  # noqa: E402

dist = pdl.distributed


def pick_device(local_rank):
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank % torch.cuda.device_count())
        return torch.device("cuda", local_rank % torch.cuda.device_count())
    return torch.device("cpu")


def backend_for(device):
    return "nccl" if device.type == "cuda" else "gloo"


def mnist_dataset(data_root=None, length=60000):
    """launch_dist.py:64-65 with this package's MNIST reader + transforms (IDX files under ``data_root``)."""
    if data_root:
        T = pdl.utils.data.transforms
        return pdl.utils.data.MNIST(root=data_root, train=True, download=True, transform=T.mnist_transform(), synthetic_length=length)
    return pdl.utils.data.SyntheticMNIST(length)


def cifar_dataset(data_root=None, length=50000):
    """example_launch.py:32-48 with this package's CIFAR10 reader + transforms (``cifar-10-batches-py`` under ``data_root``)."""
    if data_root:
        T = pdl.utils.data.transforms
        return pdl.utils.data.CIFAR10(root=data_root, train=True, download=False, transform=T.cifar10_train_transform(), synthetic_length=length)
    return pdl.utils.data.SyntheticCIFAR10(length)


def wrap(model, device):
    if device.type == "cuda":
        return nn.parallel.DistributedDataParallel(model, device_ids=[device.index], output_device=device.index)
    return nn.parallel.DistributedDataParallel(model)


def train_convnet(model, criterion, optimizer, loader, device, epochs, is_printer, graphed=False, max_steps=None,
                  checkpoint_path=None, checkpoint_every=0, start_step=0, fail_at=None):
    """Loop body of the reference's ConvNet scripts: loss printed every 100 steps by the node's local rank 0.

    Not in the reference (it has no checkpoints): every ``checkpoint_every`` steps a full training state goes to
    ``checkpoint_path`` (rank 0 writes, everyone barriers); ``start_step`` fast-forwards a resumed run past the batches
    it has already consumed; ``fail_at`` kills this process after that step (fault injection for the restart test)."""
    start = datetime.now()
    total_step = len(loader)
    if is_printer:
        print("Total step: ", total_step)
    step_fn = pdl.TrainStep(model, criterion, optimizer) if graphed else None
    seen = start_step
    for epoch in range(epochs):
        for i, (images, labels) in enumerate(loader):
            if epoch * total_step + i < start_step:
                continue  # consumed before the restart
            if step_fn is not None:
                loss = step_fn(images, labels)
            else:
                images = images.to(device, non_blocking=True)
                labels = labels.to(device, non_blocking=True)
                outputs = model(images)
                loss = criterion(outputs, labels)
                optimizer.zero_grad()
                loss.backward()
                optimizer.step()
            if (i + 1) % 100 == 0 and is_printer:
                print("Epoch [{}/{}], Step [{}/{}], Loss: {:.4f}".format(epoch + 1, epochs, i + 1, total_step, loss.item()))
            seen += 1
            if checkpoint_every and seen % checkpoint_every == 0:
                pdl.utils.checkpoint.save_checkpoint(model, checkpoint_path, optimizer=optimizer, epoch=epoch, extra={"step": seen})
            if fail_at is not None and seen == fail_at:
                os._exit(17)
            if max_steps and seen >= max_steps:
                break
        if max_steps and seen >= max_steps:
            break
    if is_printer:
        print("Training complete in: " + str(datetime.now() - start))


def train_resnet(net, criterion, optimizer, loader, device, epochs, is_printer, max_steps=None, graphed=False):
    """Loop body of the reference's ResNet scripts: running loss / accuracy every 25 steps on global rank 0."""
    if is_printer:
        print("            =======  Training  ======= \n")
    net.train()
    step_fn = pdl.TrainStep(net, criterion, optimizer) if graphed else None
    seen = 0
    metrics = pdl.utils.stats.RunningMetrics(device)  # running loss / accuracy stay on the GPU; read when a line is printed
    for ep in range(1, epochs + 1):
        metrics.reset()
        loader.sampler.set_epoch(ep)
        # batches arrive on the device one step ahead (the reference does a blocking .to(device) here, example_launch.py:83)
        for idx, (inputs, targets) in enumerate(pdl.utils.data.CudaPrefetcher(loader, device)):
            if step_fn is not None:  # the same iteration replayed as one CUDA graph (captured once per batch shape)
                loss = step_fn(inputs, targets)
                outputs = step_fn.outputs()
            else:
                outputs = net(inputs)
                loss = criterion(outputs, targets)
                optimizer.zero_grad()
                loss.backward()
                optimizer.step()
            metrics.update(outputs, targets, loss)
            if is_printer and ((idx + 1) % 25 == 0 or (idx + 1) == len(loader)):
                mean_loss, acc, _ = metrics.read()
                print("   == step: [{:3}/{}] [{}/{}] | loss: {:.3f} | acc: {:6.3f}%".format(
                    idx + 1, len(loader), ep, epochs, mean_loss, 100.0 * acc))
            seen += 1
            if max_steps and seen >= max_steps:
                break
        if max_steps and seen >= max_steps:
            break
    if is_printer:
        print("\n            =======  Training Finished  ======= \n")
