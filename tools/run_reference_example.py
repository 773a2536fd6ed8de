"""Run the reference's own example script (examples/torch_cifar10_resnet.py), UNMODIFIED, against this
repository's K-FAC through the ``kfac`` alias package - the "drops into the existing examples" check.

    python tools/run_reference_example.py [--examples DIR] [--fp16] [--iters N] [--log-dir DIR]

What the harness supplies, because this box has no network and no dataset: a synthetic stand-in for
torchvision's CIFAR10 class (random 32x32 images, same interface), and a stub for the missing
third-party ``torchsummary``.  Everything else - argument parsing, DDP, the training engine, the
optimizer / preconditioner / scheduler construction, checkpointing - is the reference's code.
The examples directory is NOT part of this repository: it is read from --examples, $KFAC_EXAMPLES,
baseline/_ref/examples (a git-ignored copy that travels to the GPU box) or /root/reference/examples.
"""
import argparse
import os
import runpy
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def find_examples(arg):
    for cand in (arg, os.environ.get('KFAC_EXAMPLES'), os.path.join(ROOT, 'baseline', '_ref', 'examples'),
                 '/root/reference/examples'):
        if cand and os.path.exists(os.path.join(cand, 'torch_cifar10_resnet.py')):
            return cand
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--examples', default=None)
    ap.add_argument('--fp16', action='store_true')
    ap.add_argument('--iters', type=int, default=6)
    ap.add_argument('--batch-size', type=int, default=64)
    ap.add_argument('--log-dir', default='/tmp/kfac_example_logs')
    ap.add_argument('--comm-method', default='comm-opt')
    args = ap.parse_args()
    examples = find_examples(args.examples)
    if examples is None:
        print('reference examples not found')
        return 77
    sys.path.insert(0, examples)
    sys.path.insert(0, ROOT)              # `import kfac` -> the alias package of this repository

    import numpy as np
    import torch
    import torchvision
    from PIL import Image

    import kfac
    assert os.path.realpath(kfac.__file__).startswith(os.path.realpath(ROOT)), kfac.__file__

    class SyntheticCIFAR10(torch.utils.data.Dataset):
        """Same constructor and item type (PIL image, int label) as torchvision.datasets.CIFAR10."""

        def __init__(self, root, train=True, download=False, transform=None):
            n = args.iters * args.batch_size if train else 2 * args.batch_size
            rng = np.random.RandomState(0 if train else 1)
            self.data = rng.randint(0, 256, size=(n, 32, 32, 3), dtype=np.uint8)
            self.targets = rng.randint(0, 10, size=n).tolist()
            self.transform = transform

        def __len__(self):
            return len(self.targets)

        def __getitem__(self, i):
            img = Image.fromarray(self.data[i])
            if self.transform is not None:
                img = self.transform(img)
            return img, self.targets[i]

    torchvision.datasets.CIFAR10 = SyntheticCIFAR10
    stub = types.ModuleType('torchsummary')
    stub.summary = lambda *a, **k: None
    sys.modules.setdefault('torchsummary', stub)

    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29931')
    os.environ.setdefault('RANK', '0')
    os.environ.setdefault('WORLD_SIZE', '1')
    os.environ.setdefault('LOCAL_RANK', '0')
    sys.argv = ['torch_cifar10_resnet.py', '--epochs', '1', '--batch-size', str(args.batch_size),
                '--val-batch-size', str(args.batch_size), '--kfac-update-freq', '2', '--kfac-cov-update-freq', '1',
                '--log-dir', args.log_dir, '--checkpoint-freq', '1', '--kfac-comm-method', args.comm_method,
                '--data-dir', os.path.join(args.log_dir, 'data')] + (['--fp16'] if args.fp16 else [])
    os.chdir(examples)
    runpy.run_path(os.path.join(examples, 'torch_cifar10_resnet.py'), run_name='__main__')
    ckpt = torch.load(os.path.join(args.log_dir, 'checkpoint_1.pth.tar'), weights_only=False)
    layers = ckpt['preconditioner']['layers']
    print('EXAMPLE OK: %d K-FAC layers in the checkpoint, step %d, A[0] %s %s'
          % (len(layers), ckpt['preconditioner']['param_groups'][0]['step'], tuple(layers[0]['A'].shape),
             layers[0]['A'].dtype))
    torch.distributed.destroy_process_group()
    return 0


if __name__ == '__main__':
    sys.exit(main())
