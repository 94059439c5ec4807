import importlib, sys, torch
sys.path.insert(0, "/root/repo")
pkg = importlib.import_module("akarirender-1_b200")
defaults = {"finish_max": 65536, "finish_div": 4, "check_every": 4, "tail_share": 1, "refill_below": 0, "tail_share_min": 16}
for wl in ("c3", "c5", "c1"):
    if wl == "c3":
        sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=64, max_depth=8, materials="c3")
    elif wl == "c5":
        sc = pkg.scenes.many_light_scene(n=708, resolution=(1920, 1080), spp=32, max_depth=16, n_quads=5000, rr_start_depth=3)
    else:
        sc = pkg.scenes.cornell_box((512, 512), 16, 5)
    W, H = sc.camera.resolution
    be = pkg.Backend(sc, 0)
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    for opts in ({}, {"tail_share_min": 24}, {"tail_share_min": 16}, {"tail_share_min": 12}, {"tail_share_min": 8}, {"tail_share_min": 4}, {"tail_share": 2}, {}):
        for k, v in opts.items(): pkg.set_option(k, v)
        best = min(be.render_device(film.data_ptr(), p).ms_total for _ in range(4))
        print(wl, opts, "%.3f ms" % best, flush=True)
        for k in opts: pkg.set_option(k, defaults[k])
    be.close()
