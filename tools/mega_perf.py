import importlib, sys, torch
sys.path.insert(0, "/root/repo")
pkg = importlib.import_module("akarirender-1_b200")
sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=16, max_depth=8, materials="c3")
be = pkg.Backend(sc, 0)
film = torch.zeros(1080 * 1920 * 4, dtype=torch.float32, device="cuda")
for world, rank in ((1, 0), (8, 0)):
    p = sc.integrator.params(rank=rank, world_size=world, shard_tile=32)
    p.mode = pkg.abi.MODE_MEGAKERNEL
    best = min(be.render_device(film.data_ptr(), p).ms_total for _ in range(3))
    print("megakernel C3 16 spp share 1/%d: %.2f ms" % (world, best), flush=True)
