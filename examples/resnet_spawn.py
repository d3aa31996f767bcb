"""ResNet-18 / CIFAR-10 with multiprocessing.spawn and tcp:// init (counterpart of the reference's example_mp.py).

    python examples/resnet_spawn.py --nodes 1 --ngpus_per_node 2 --dist-url tcp://127.0.0.1:12345 --node_rank 0
"""
import argparse

import torch

from _common import cifar_dataset, dist, nn, pdl, pick_device, train_resnet, wrap

BATCH_SIZE = 256
EPOCHS = 5

parser = argparse.ArgumentParser()
parser.add_argument("--nodes", default=1, type=int, help="number of nodes for distributed training")
parser.add_argument("--ngpus_per_node", default=2, type=int, help="number of GPUs per node for distributed training")
parser.add_argument("--dist-url", default="tcp://127.0.0.1:12345", type=str, help="url used to set up distributed training")
parser.add_argument("--node_rank", default=0, type=int, help="node rank for distributed training")
parser.add_argument("--data-root", default=None)
parser.add_argument("--max-steps", type=int, default=None)
parser.add_argument("--graphed", action="store_true", help="run each iteration as one CUDA graph (pdl.TrainStep)")


def train_worker(local_rank, ngpus_per_node, args):
    args.global_rank = args.node_rank * ngpus_per_node + local_rank
    dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo", init_method=args.dist_url,
                            world_size=args.global_world_size, rank=args.global_rank)
    device = pick_device(local_rank)
    print(f"[init] == local rank: {local_rank}, global rank: {args.global_rank} ==")
    net = pdl.models.resnet18(pretrained=False, num_classes=10).to(device)
    net = wrap(net, device)
    train_data = cifar_dataset(args.data_root)
    sampler = pdl.utils.data.distributed.DistributedSampler(train_data, shuffle=True)
    loader = torch.utils.data.DataLoader(train_data, batch_size=BATCH_SIZE // max(1, ngpus_per_node), num_workers=0,
                                         pin_memory=device.type == "cuda", sampler=sampler)
    criterion = nn.CrossEntropyLoss().to(device)
    optimizer = pdl.optim.SGD(net.parameters(), lr=0.01 * 2, momentum=0.9, weight_decay=0.0001, nesterov=True)
    train_resnet(net, criterion, optimizer, loader, device, EPOCHS, is_printer=(args.global_rank == 0), max_steps=args.max_steps,
                 graphed=args.graphed)
    dist.destroy_process_group()


def main():
    args = parser.parse_args()
    args.global_world_size = args.ngpus_per_node * args.nodes
    pdl.multiprocessing.spawn(train_worker, nprocs=args.ngpus_per_node, args=(args.ngpus_per_node, args))


if __name__ == "__main__":
    main()
