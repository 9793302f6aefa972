"""units.dataloader — drop-in for the reference's units/dataloader.py (called by main.py:32-38): the same three
functions with the same arguments and the same torchvision pipeline, so the reference's main.py runs unchanged on this
tree.  Host-side input pipeline only: nothing here touches the GPU; the training step receives
the batches exactly as it does in the reference.

  get_dataset(dataroot, size=64)   ImageFolder -> ToTensor -> RandomHorizontalFlip -> bicubic Resize((size, size)) ->
                                   Normalize(ImageNet mean / std)                                  (reference :5-15)
  get_MNIST(dataroot)              MNIST train split -> ToTensor -> bicubic Resize((64, 64)) -> Normalize(0.5, 0.5),
                                   download=True                                                   (reference :16-27)
  get_dataloader(dataset, batch)   DataLoader(shuffle=True, drop_last=True), single process        (reference :29-32)
"""
import torch
import torchvision.datasets as dset
import torchvision.transforms as transforms

_IMAGENET_MEAN, _IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def _pipeline(size, mean, std, flip):
    steps = [transforms.ToTensor()]
    if flip:
        steps.append(transforms.RandomHorizontalFlip())
    steps += [transforms.Resize((size, size), interpolation=transforms.InterpolationMode.BICUBIC),
              transforms.Normalize(mean, std)]
    return transforms.Compose(steps)


def get_dataset(dataroot, size=64):
    return dset.ImageFolder(root=dataroot, transform=_pipeline(size, _IMAGENET_MEAN, _IMAGENET_STD, flip=True))


def get_MNIST(dataroot):
    return dset.MNIST(root=dataroot, train=True, transform=_pipeline(64, (0.5), (0.5), flip=False), download=True)


def get_dataloader(dataset, batch_size):
    return torch.utils.data.DataLoader(dataset, batch_size=batch_size, shuffle=True, drop_last=True)
