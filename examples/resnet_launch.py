"""ResNet-18 / CIFAR-10 with the launcher (counterpart of the reference's example_launch.py).

    python -m pytorch_distributed_learning_b200.launch --nproc_per_node=2 --use_env examples/resnet_launch.py
"""
import argparse
import os

import torch

from _common import backend_for, cifar_dataset, dist, nn, pdl, pick_device, train_resnet, wrap

BATCH_SIZE = 128
EPOCHS = 5

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--local-rank", "--local_rank", type=int, default=None)
    ap.add_argument("--data-root", default=None)
    ap.add_argument("--max-steps", type=int, default=None)
    ap.add_argument("--graphed", action="store_true", help="run each iteration as one CUDA graph (pdl.TrainStep)")
    args = ap.parse_args()
    rank, local_rank = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    device = pick_device(rank)
    dist.init_process_group(backend=backend_for(device))
    print(f"[init] == local rank: {local_rank}, global rank: {rank} ==")
    net = pdl.models.resnet18(pretrained=False, num_classes=10).to(device)
    net = wrap(net, device)            # wrap first, optimizer afterwards -- as the reference does in this script
    train_data = cifar_dataset(args.data_root)
    sampler = pdl.utils.data.distributed.DistributedSampler(train_data, shuffle=True)
    loader = torch.utils.data.DataLoader(train_data, batch_size=BATCH_SIZE, num_workers=0 if args.data_root is None else 4,
                                         pin_memory=device.type == "cuda", sampler=sampler)
    criterion = nn.CrossEntropyLoss()
    optimizer = pdl.optim.SGD(net.parameters(), lr=0.01 * 2, momentum=0.9, weight_decay=0.0001, nesterov=True)
    train_resnet(net, criterion, optimizer, loader, device, EPOCHS, is_printer=(rank == 0), max_steps=args.max_steps, graphed=args.graphed)
    dist.destroy_process_group()
