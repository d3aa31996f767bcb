"""ConvNet / MNIST with the launcher (counterpart of the reference's launch_dist.py).

    python -m pytorch_distributed_learning_b200.launch --nproc_per_node=8 --use_env examples/convnet_launch.py [--graphed]
"""
import argparse
import os

import torch

from _common import backend_for, dist, mnist_dataset, nn, pdl, pick_device, train_convnet, wrap

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--local-rank", "--local_rank", type=int, default=None)  # appended by the launcher unless --use_env
    ap.add_argument("--epochs", type=int, default=10)
    ap.add_argument("--data-root", default=None)
    ap.add_argument("--graphed", action="store_true", help="run each iteration as one CUDA graph (pdl.TrainStep)")
    ap.add_argument("--max-steps", type=int, default=None)
    ap.add_argument("--save", default=None, help="write a rank-0 state_dict checkpoint here at the end")
    # beyond the reference (which never checkpoints): periodic training-state checkpoints + resume, for `launch --max_restarts`
    ap.add_argument("--checkpoint", default=None, help="training-state checkpoint: written every --checkpoint-every steps, resumed from if present")
    ap.add_argument("--checkpoint-every", type=int, default=0)
    ap.add_argument("--inject-failure-at", type=int, default=None, help="fault injection: rank 1 dies after this step of the FIRST attempt")
    args = ap.parse_args()

    rank, local_rank = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    device = pick_device(rank)
    dist.init_process_group(backend=backend_for(device))
    torch.manual_seed(0)
    model = pdl.models.ConvNet().to(device)
    batch_size = 100
    criterion = nn.CrossEntropyLoss().to(device)
    optimizer = pdl.optim.SGD(model.parameters(), 1e-4)   # optimizer first, then the wrap -- as the reference does
    model = wrap(model, device)

    train_data = mnist_dataset(args.data_root)
    sampler = pdl.utils.data.distributed.DistributedSampler(train_data, num_replicas=dist.get_world_size(), rank=local_rank)
    loader = torch.utils.data.DataLoader(dataset=train_data, batch_size=batch_size, shuffle=False, pin_memory=device.type == "cuda", sampler=sampler)
    start_step = 0
    if args.checkpoint and os.path.exists(args.checkpoint):
        info = pdl.utils.checkpoint.load_checkpoint(model, args.checkpoint, optimizer=optimizer)
        start_step = int(info.get("extra", {}).get("step", 0))
        if local_rank == 0:
            print("Resumed from %s at step %d (restart %s)" % (args.checkpoint, start_step, os.environ.get("TORCHELASTIC_RESTART_COUNT", "0")))
    first_attempt = os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") == "0"
    train_convnet(model, criterion, optimizer, loader, device, args.epochs, is_printer=(local_rank == 0), graphed=args.graphed, max_steps=args.max_steps,
                  checkpoint_path=args.checkpoint, checkpoint_every=args.checkpoint_every if args.checkpoint else 0, start_step=start_step,
                  fail_at=args.inject_failure_at if (first_attempt and rank == 1) else None)
    if args.save:
        pdl.utils.checkpoint.save_checkpoint(model, args.save)
    dist.destroy_process_group()
