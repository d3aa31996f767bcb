"""ConvNet / MNIST with multiprocessing.spawn (counterpart of the reference's mpspawn_dist.py).

    python examples/convnet_spawn.py -n 1 -g 8 -nr 0 --epochs 2
"""
import argparse
import os

import torch

from _common import backend_for, dist, mnist_dataset, nn, pdl, pick_device, train_convnet, wrap


def train(gpu, args):
    print("My rank is", args.nr * args.gpus + gpu)
    rank = args.nr * args.gpus + gpu
    dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo", init_method="env://", world_size=args.world_size, rank=rank)
    torch.manual_seed(0)
    model = pdl.models.ConvNet()
    device = pick_device(gpu)          # device is bound AFTER init here, as in the reference's spawn script
    model = model.to(device)
    print("load model sucessfully!")
    batch_size = 100
    criterion = nn.CrossEntropyLoss().to(device)
    optimizer = pdl.optim.SGD(model.parameters(), 1e-4)
    model = wrap(model, device)
    print("wrap model sucessfully!")
    train_data = mnist_dataset(args.data_root)
    sampler = pdl.utils.data.distributed.DistributedSampler(train_data, num_replicas=args.world_size, rank=rank)
    loader = torch.utils.data.DataLoader(dataset=train_data, batch_size=batch_size, shuffle=False, num_workers=0,
                                         pin_memory=device.type == "cuda", sampler=sampler)
    print("load data sucessfully!")
    train_convnet(model, criterion, optimizer, loader, device, args.epochs, is_printer=(gpu == 0), graphed=args.graphed, max_steps=args.max_steps)
    dist.destroy_process_group()


def main():
    parser = argparse.ArgumentParser()
    parser.add_argument("-n", "--nodes", default=1, type=int, metavar="N", help="number of nodes")
    parser.add_argument("-g", "--gpus", default=1, type=int, help="number of processes (GPUs) per node")
    parser.add_argument("-nr", "--nr", default=0, type=int, help="ranking of this node within the nodes")
    parser.add_argument("--epochs", default=2, type=int, metavar="N", help="number of total epochs to run")
    parser.add_argument("--master-addr", default="127.0.0.1", help="the reference hard-codes 172.16.16.5 here")
    parser.add_argument("--master-port", default="8888")
    parser.add_argument("--data-root", default=None)
    parser.add_argument("--graphed", action="store_true")
    parser.add_argument("--max-steps", type=int, default=None)
    args = parser.parse_args()
    args.world_size = args.gpus * args.nodes
    os.environ["MASTER_ADDR"] = args.master_addr
    os.environ["MASTER_PORT"] = args.master_port
    pdl.multiprocessing.spawn(train, nprocs=args.gpus, args=(args,))


if __name__ == "__main__":
    main()
