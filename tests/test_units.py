"""units/mixup.py and units/Weight_Initialization.py drop-ins against the unmodified reference (when mounted) and against
the plain formula; the Beta draws must be bit-identical (host numpy RNG)."""
import numpy as np
import pytest
import torch

from conftest import needs_reference


def _both(fn_name, true, fake, alpha, ref_mod, my_mod):
    np.random.seed(3)
    want = getattr(ref_mod, fn_name)(true, fake, alpha, true.device)
    np.random.seed(3)
    got = getattr(my_mod, fn_name)(true, fake, alpha, true.device)
    return got, want


@needs_reference
def test_mixup_matches_reference_bit_exact_coefficients(ref_backend):
    from oracle import reference as R
    ref = R.load("units.mixup")
    from units import mixup as mine
    t, f = torch.randn(64, 3, 64, 64), torch.randn(64, 3, 64, 64)
    got, want = _both("mixup_data", t, f, 0.4, ref, mine)
    assert got[2] == want[2]                                   # lam: bit exact
    assert torch.equal(got[1], want[1])
    assert torch.allclose(got[0], want[0], atol=1e-6)
    got, want = _both("mixup_data_wgan", t, f, 0.4, ref, mine)
    assert torch.equal(got[1], want[1])                        # per-row Beta labels: bit exact, float64
    assert got[0].dtype == want[0].dtype == torch.float64
    assert torch.allclose(got[0], want[0], atol=1e-12)
    for alpha in (0, -1.0):
        got, want = _both("mixup_data", t, f, alpha, ref, mine)
        assert got[2] == want[2] == 1
    # the (B,1) x (B,3,64,64) broadcast only works for B in {1, 64}: same exception type otherwise
    t8, f8 = torch.randn(8, 3, 64, 64), torch.randn(8, 3, 64, 64)
    with pytest.raises(RuntimeError):
        ref.mixup_data_wgan(t8, f8, 0.4, t8.device)
    with pytest.raises(RuntimeError):
        mine.mixup_data_wgan(t8, f8, 0.4, t8.device)


def test_weights_init_by_class_name():
    import torch.nn as nn
    from units.Weight_Initialization import weights_init
    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv2d(3, 4, 3), nn.BatchNorm2d(4), nn.Linear(4, 4))
    before = net[2].weight.clone()
    net.apply(weights_init)
    assert abs(float(net[0].weight.std()) - 0.02) < 0.01
    assert abs(float(net[1].weight.mean()) - 1.0) < 0.05 and float(net[1].bias.abs().sum()) == 0
    assert torch.equal(net[2].weight, before)


@pytest.mark.gpu
def test_mixup_on_gpu():
    from units import mixup as mine
    t, f = torch.randn(64, 3, 64, 64).cuda(), torch.randn(64, 3, 64, 64).cuda()
    np.random.seed(5)
    mixed, label, lam = mine.mixup_data(t, f, 0.4, t.device)
    np.random.seed(5)
    assert lam == np.random.beta(0.4, 0.4)
    assert torch.allclose(mixed, lam * t + (1 - lam) * f, atol=1e-6)
    np.random.seed(5)
    mixed, label = mine.mixup_data_wgan(t, f, 0.4, t.device)
    np.random.seed(5)
    lab = torch.from_numpy(np.random.beta(0.4, 0.4, size=(64, 1))).cuda()
    assert torch.equal(label, lab) and mixed.dtype == torch.float64
    assert torch.allclose(mixed, lab * t + (1 - lab) * f, atol=1e-12)


def test_factories_mirror_the_reference_names():
    """units/get_generators.py / get_discriminator.py: every reference factory name exists with the same argument
    list; carried families build, the others say they are out of scope."""
    import inspect
    from units import get_discriminator, get_generators
    cpu = torch.device("cpu")
    g = get_generators.get_1(cpu, 256, (3, 64, 64))
    d = get_discriminator.get_1(cpu, (3, 64, 64))
    assert sum(p.numel() for p in g.parameters()) == 256 * 256 + 256 + 256 * 512 + 512 + 512 * 12288 + 12288
    assert sum(p.numel() for p in d.parameters()) == 12288 * 256 + 256 + 256 * 64 + 64 + 64 + 1
    assert list(inspect.signature(get_generators.get_3_progan).parameters) == ["ngpu", "device", "nz", "ngf", "nc"]
    assert list(inspect.signature(get_generators.get_13).parameters) == ["ngpu", "device", "z_dim"]
    assert inspect.signature(get_generators.get_13).parameters["z_dim"].default == 128
    assert list(inspect.signature(get_discriminator.get_3_wgan_progan).parameters) == ["ngpu", "device", "ndf", "nc"]
    assert list(inspect.signature(get_discriminator.get_7_wgansn).parameters) == ["ngpu", "device"]
    for n in ("2", "4", "5", "6", "7", "8", "9", "10", "10_2", "11", "12", "13"):
        assert callable(getattr(get_generators, "get_" + n))
    for n in ("2", "3_wgan", "3_wgansn", "4", "5", "5_wgan", "5_wgansn", "6_wgan", "7_wgan", "8", "9"):
        assert callable(getattr(get_discriminator, "get_" + n))
    with pytest.raises(NotImplementedError):       # every model file of the reference has a drop-in; anything else says so
        get_generators._module("generator_99")
    g4 = get_generators.get_4(1, cpu, z_dim=32)
    assert type(g4).__module__ == "generators.generator_4" and len(g4.model) == 14
    d6 = get_discriminator.get_6_wgan(1, cpu)  # every discriminator factory of the reference builds here
    assert type(d6).__module__ == "discriminators.discriminator_6_wgangp" and len(d6.model) == 8
    d9 = get_discriminator.get_9(1, cpu)       # the self-attention critic: 64 base features, five attention layers
    assert type(d9).__name__ == "Discriminator" and sum(type(m).__name__ == "SelfAttention" for m in d9.conv) == 5
    d4 = get_discriminator.get_3_wgan_progan(1, cpu, ndf=8, nc=3)
    assert type(d4).__name__ == "Discriminator" and len(d4.main) == 24
    # get_2 (both factories) applies the DCGAN weights_init: conv weights N(0, 0.02), BatchNorm weights N(1, 0.02)
    torch.manual_seed(0)
    g2, d2 = get_generators.get_2(1, cpu, 16, 8, 3), get_discriminator.get_2(1, cpu, 8, 3)
    assert abs(float(g2.main[0].weight.std()) - 0.02) < 0.005 and abs(float(g2.main[1].weight.mean()) - 1.0) < 0.02
    assert abs(float(d2.main[2].weight.std()) - 0.02) < 0.005 and float(d2.main[3].bias.abs().max()) == 0.0
    # the other carried critics build through their factories with the reference's widths
    count = lambda m: sum(p.numel() for p in m.parameters())
    assert count(get_discriminator.get_3_wgansn(1, cpu, 8, 3)) == 3 * 8 * 16 + 16 * (8 * 16 + 16 * 32 + 32 * 64) + 64 * 16
    assert count(get_discriminator.get_7_wgan(1, cpu)) == 2747849
    assert count(get_discriminator.get_5_wgan(1, cpu)) == 25338113
    assert count(get_discriminator.get_5_wgansn(1, cpu)) == 26103041
    assert count(get_discriminator.get_5(1, cpu)) == 26118657


def test_dataloader_mirrors_the_reference_pipeline(tmp_path):
    """units/dataloader.py: same three functions as the reference's (units/dataloader.py:5-32) — ImageFolder ->
    ToTensor -> flip -> bicubic Resize((size, size)) -> ImageNet Normalize; DataLoader(shuffle, drop_last)."""
    import numpy as np
    from PIL import Image
    from units import dataloader
    root = tmp_path / "imgs" / "cls"
    root.mkdir(parents=True)
    rs = np.random.RandomState(0)
    for i in range(5):
        Image.fromarray(rs.randint(0, 255, (40 + i, 52, 3), dtype=np.uint8)).save(root / f"{i}.png")
    ds = dataloader.get_dataset(str(tmp_path / "imgs"), size=64)
    assert len(ds) == 5
    x, y = ds[0]
    assert x.shape == (3, 64, 64) and x.dtype == torch.float32 and y == 0
    names = [type(t).__name__ for t in ds.transform.transforms]
    assert names == ["ToTensor", "RandomHorizontalFlip", "Resize", "Normalize"]
    assert tuple(ds.transform.transforms[3].mean) == (0.485, 0.456, 0.406)
    dl = dataloader.get_dataloader(ds, 2)
    batches = list(dl)
    assert len(dl) == 2 and len(batches) == 2 and batches[0][0].shape == (2, 3, 64, 64)   # drop_last
    assert dl.drop_last and type(dl.sampler).__name__ == "RandomSampler"
    assert callable(dataloader.get_MNIST)


def test_cached_loader_equals_the_dataset_pipeline_and_shards_quarter_major(tmp_path):
    """gan_b200.data.CachedImageLoader: every yielded image is the dataset's own pipeline output
    (un-flipped, or flipped left-right), an epoch visits each kept sample once, drop_last, and two ranks together hold
    exactly the single-process global batch in quarter-major order."""
    import numpy as np
    from PIL import Image
    from gan_b200.data import CachedImageLoader
    from gan_b200.dist import shard_quarter_major
    from units import dataloader
    root = tmp_path / "imgs" / "cls"
    root.mkdir(parents=True)
    rs = np.random.RandomState(1)
    for i in range(19):
        Image.fromarray(rs.randint(0, 255, (33 + i, 47, 3), dtype=np.uint8)).save(root / f"{i:02d}.png")
    ds = dataloader.get_dataset(str(tmp_path / "imgs"), size=64)
    single = CachedImageLoader(ds, 8, seed=3)
    assert [type(t).__name__ for t in ds.transform.transforms][1] == "RandomHorizontalFlip"  # the dataset is left as it was
    assert single.flip_p == 0.5 and len(single) == 2 and single.images.shape == (19, 3, 64, 64)
    plain = [ds.transform.transforms[3](ds.transform.transforms[2](ds.transform.transforms[0](Image.open(p).convert("RGB"))))
             for p, _ in ds.samples]
    for i in range(19):
        assert torch.equal(single.images[i], plain[i])
    seen, n_flipped = [], 0
    for x, y in single:
        assert x.shape == (8, 3, 64, 64) and y.shape == (8,) and int(y.max()) == 0
        for img in x:
            hit = [i for i in range(19) if torch.equal(img, plain[i]) or torch.equal(img, plain[i].flip(-1))]
            assert len(hit) == 1
            seen.append(hit[0])
            n_flipped += int(not torch.equal(img, plain[hit[0]]))
    assert len(seen) == 16 == len(set(seen)) and 0 < n_flipped < 16
    second = torch.cat([x for x, _ in single])  # a second epoch: new permutation / flips (seed + epoch)
    first = torch.cat([x for x, _ in CachedImageLoader(ds, 8, seed=3)])
    assert second.shape == first.shape and not torch.equal(second, first)
    # two ranks: same permutation and flips, quarter-major slices of each global batch of 8
    g = CachedImageLoader(ds, 8, seed=3)
    r0, r1 = (CachedImageLoader(ds, 4, rank=r, world=2, seed=3) for r in (0, 1))
    for (xg, _), (x0, _), (x1, _) in zip(g, r0, r1):
        assert torch.equal(x0, shard_quarter_major(xg, 0, 2)) and torch.equal(x1, shard_quarter_major(xg, 1, 2))
        # the ranks' MiniBatchStdDev rows (x.view(4, -1)) side by side are the global batch's rows
        assert torch.equal(torch.cat([x0.reshape(4, -1), x1.reshape(4, -1)], 1), xg.reshape(4, -1))
    with pytest.raises(ValueError):
        CachedImageLoader(ds, 6, rank=0, world=2)


@pytest.mark.gpu
def test_device_cached_loader_equals_the_host_path_bit_for_bit(tmp_path):
    """CachedImageLoader(device_cache=True): the data set lives in HBM and every batch is ONE ganb200_gather_flip launch;
    same permutation, same flips, same bytes as the host-gather path (and therefore as the dataset's own pipeline)."""
    import numpy as np
    from PIL import Image
    from gan_b200.data import CachedImageLoader
    from units import dataloader
    root = tmp_path / "imgs" / "cls"
    root.mkdir(parents=True)
    rs = np.random.RandomState(2)
    for i in range(37):
        Image.fromarray(rs.randint(0, 255, (40 + i, 51, 3), dtype=np.uint8)).save(root / f"{i:02d}.png")
    ds = dataloader.get_dataset(str(tmp_path / "imgs"), size=64)
    host = CachedImageLoader(ds, 8, seed=5)
    dev = CachedImageLoader(ds, 8, device="cuda", seed=5, device_cache=True)
    n = 0
    for _ in range(2):  # two epochs: permutation and flips change with seed + epoch on both
        for (xh, yh), (xd, yd) in zip(host, dev):
            assert xd.is_cuda and torch.equal(xd.cpu(), xh) and torch.equal(yd.cpu(), yh)
            n += 1
    assert n == 2 * len(host) == 8
