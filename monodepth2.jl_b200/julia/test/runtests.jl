# Test suite for the CuArray{Float32} methods that MDB200.jl adds to Monodepth (libmdb200.so behind `ccall`).
# It asserts the same facts as the reference's own suite (test/runtests.jl testsets "Test rotations",
# "Test transformation", "Test SSIM", "Test smooth loss", "Test disparity -> depth", "Test identity image warping",
# "Test pose derivative"), but on device arrays, and then holds every GPU method against the reference's pure-Julia
# CPU method of the same function on the same random inputs.  Written for this repository (not a copy of the reference
# file).  NOT EXECUTED on the build / GPU boxes (no Julia runtime, SURVEY F2); the Python harness tests/test_gpu_runtests.py
# runs the same scenarios through the same C ABI.
#
# Run from a checkout of jumerckx/Monodepth2.jl with `include("MDB200.jl")` added to src/Monodepth.jl:
#   julia --project -e 'include("path/to/monodepth2.jl_b200/julia/test/runtests.jl")'
using Test, CUDA, Zygote, LinearAlgebra, Random, Statistics
import Flux, NNlib
using FixedPointNumbers: N0f8
using ImageCore: Gray, channelview
using ImageTransformations: imresize
using Monodepth
const M = Monodepth
CUDA.allowscalar(false)
Random.seed!(7)

rotvec_matrix(v) = (θ = norm(v); k = v ./ θ; Kx = [0 -k[3] k[2]; k[3] 0 -k[1]; -k[2] k[1] 0]; I + sin(θ) * Kx + (1 - cos(θ)) * Kx^2)

@testset "MDB200 (CuArray{Float32} methods)" begin

@testset "so3 / hat / composeT" begin
    v = Float32[0.3, -0.2, 0.5]
    R = Array(M.so3_exp_map(cu(reshape(v, 3, 1))))[:, :, 1]
    @test R ≈ rotvec_matrix(Float64.(v)) atol = 1e-5
    S = Array(M.hat(cu(reshape(v, 3, 1))))[:, :, 1]
    @test S ≈ Float32[0 -v[3] v[2]; v[3] 0 -v[1]; -v[2] v[1] 0]
    # pullback of hat: same numbers as the reference's rrule on the CPU
    d = rand(Float32, 3, 3, 1)
    g_gpu = Array(Zygote.gradient(r -> sum(M.hat(r) .* cu(d)), cu(reshape(v, 3, 1)))[1])
    g_cpu = Zygote.gradient(r -> sum(M.hat(r) .* d), reshape(v, 3, 1))[1]
    @test g_gpu ≈ g_cpu atol = 1e-6
    # composeT forward, and invert = true undoing it
    rv, tv = cu(reshape(v, 3, 1)), cu(reshape(Float32[0.1, 0.2, -0.3], 3, 1, 1))
    p = rand(Float32, 3, 1, 1)
    Rf, tf = M.composeT(rv, tv, false)
    q = Array(Rf)[:, :, 1] * p[:, :, 1] .+ Array(tf)[:, :, 1]
    @test q ≈ rotvec_matrix(Float64.(v)) * p[:, :, 1] .+ Array(tv)[:, :, 1] atol = 1e-6
    Ri, ti = M.composeT(rv, tv, true)
    @test Array(Ri)[:, :, 1] * q .+ Array(ti)[:, :, 1] ≈ p[:, :, 1] atol = 1e-5
end

@testset "SSIM" begin
    ssim = M.SSIM()
    one_, zero_ = CUDA.ones(Float32, 2, 2, 1, 2), CUDA.zeros(Float32, 2, 2, 1, 2)
    @test all(Array(ssim(one_, one_)) .== 0)
    @test all(isapprox.(Array(ssim(one_, zero_)), 0.5f0; atol = 0.1f0))
    a, b = CUDA.rand(Float32, 9, 7, 3, 2), CUDA.rand(Float32, 9, 7, 3, 2)
    @test Array(ssim(a, b)) ≈ Array(ssim(b, a)) atol = 1e-6
    @test Array(ssim(a, b)) ≈ M.SSIM()(Array(a), Array(b)) atol = 2e-6          # GPU method vs the reference's CPU method
end

@testset "smooth loss" begin
    disp = reshape(Float32[0 0.1; 0.2 0.3], 2, 2, 1)
    img = reshape(Float32[0.1 0.2; 0.3 0.4], 2, 2, 1, 1)
    @test M.smooth_loss(cu(disp), cu(img)) ≈ 0.2542f0 atol = 1e-4
    d, im = rand(Float32, 12, 9, 2), rand(Float32, 12, 9, 3, 2)
    @test M.smooth_loss(cu(d), cu(im)) ≈ M.smooth_loss(d, im) rtol = 1e-5
    g_gpu = Array(Zygote.gradient(z -> M.smooth_loss(z, cu(im)), cu(d))[1])
    g_cpu = Zygote.gradient(z -> M.smooth_loss(z, im), d)[1]
    @test g_gpu ≈ g_cpu atol = 1e-4 * maximum(abs, g_cpu)
end

@testset "disparity -> depth" begin
    d = Array(M.disparity_to_depth(CUDA.rand(Float32, 5, 4, 1, 2), 0.1, 100.0))
    @test all(0.1f0 - 1f-6 .<= d .<= 100f0 + 1f-3)
end

@testset "identity warp" begin
    res, N = 16, 2
    K = Float32[910 0 res/2; 0 910 res/2; 0 0 1]; invK = Float32.(Base.inv(Float64.(K)))
    img = CUDA.rand(Float32, res, res, 1, N)
    x = repeat(reshape(img, res, res, 1, 1, N), 1, 1, 1, 3, 1)     # (W,H,C,L=3,N): three identical frames
    disp = CUDA.rand(Float32, 1, res, res, N)
    Ps = [M.composeT(CUDA.zeros(Float32, 3, N) .+ 1f-6, CUDA.zeros(Float32, 3, 1, N), false)]
    w = M.warp(disp, x, Ps, nothing, nothing, invK, K; min_depth = 0.1, max_depth = 100.0, source_ids = [1])
    @test Array(w[1]) ≈ Array(img) atol = 1e-3
end

@testset "slow_depth closure, function by function vs fused" begin
    W, H, N = 48, 20, 1
    K = Float32[66 0 W/2; 0 66 H/2; 0 0 1]; invK = Float32.(Base.inv(Float64.(K)))
    x = CUDA.rand(Float32, W, H, 3, 3, N)
    disp = CUDA.rand(Float32, 1, W, H, N) .* 0.5f0 .+ 0.2f0
    poses = [M.Pose(cu(0.01f0 .* randn(Float32, 3, N)), cu(0.05f0 .* randn(Float32, 3, 1, N))) for _ in 1:2]
    ssim = M.SSIM(); target = x[:, :, :, 2, :]
    function closure(disp, poses)
        Ps = [M.composeT(p.rvec, p.tvec, sid < 2) for (p, sid) in zip(poses, [1, 3])]
        warped = M.warp(disp, x, Ps, nothing, nothing, invK, K; min_depth = 0.1, max_depth = 100.0, source_ids = [1, 3])
        sum(M.prediction_loss(ssim, warped, target)) / (W * H * N) + M.smooth_loss(reshape(disp, W, H, N), target)
    end
    l, back = Zygote.pullback(closure, disp, poses)
    gd, gp = back(1f0)
    cache = M.TrainCache(ssim, nothing, nothing, K, invK, 2, [1, 3], [1.0])
    params = M.Params(min_depth = 0.1, max_depth = 100.0, automasking = false, target_size = (W, H), batch_size = N)
    lf, dd, dp = M.MDB200.fused_loss(x, [reshape(disp, W, H, 1, N)], poses, nothing, cache, params;
                                     normalize_disparity = false, smooth_weight_mode = 1)
    @test l ≈ Array(lf)[1] rtol = 1e-5
    @test Array(gd) ≈ reshape(Array(dd[1]), 1, W, H, N) atol = 1e-4 * maximum(abs, Array(dd[1]))
    @test Array(gp[1].rvec) ≈ Array(dp[1].rvec) atol = 1e-4 * maximum(abs, Array(dp[1].rvec))
    # the same step from HOST arrays through the one C call that pipelines copy-in, loss and copy-out
    plan = M.MDB200.HostPlan(W, H, 3, N, [(W, H)], cache, params; normalize_disparity = false, smooth_weight_mode = 1)
    hposes = [(rvec = Array(p.rvec), tvec = Array(p.tvec)) for p in poses]
    lh, ddh, dph = M.MDB200.fused_loss_host(plan, Array(x), [reshape(Array(disp), W, H, 1, N)], hposes, nothing)
    @test lh ≈ Array(lf)[1] rtol = 1e-6
    @test ddh[1] ≈ Array(dd[1]) rtol = 1e-5
    @test dph[2].tvec ≈ Array(dp[2].tvec) rtol = 1e-5
end

@testset "render operators" begin
    W, H, Np, B = 24, 16, 4, 2
    K = Float32[30 0 W/2; 0 30 H/2; 0 0 1]; K_inv = Float32.(Base.inv(Float64.(K)))
    @test Array(M.inv(cu(K))) ≈ K_inv atol = 1e-5
    mesh = M.MDB200.create_meshgrid_gpu(H, W)
    @test Array(mesh) ≈ Float32.(M.create_meshgrid(H, W))
    disparity = cu(Float32.(repeat(range(1, 0.1; length = Np), 1, B)))
    xyz_src = M.get_src_xyz_from_plane_disparity(mesh, disparity, K_inv)
    pose = M.Pose(cu(0.01f0 .* randn(Float32, 3, B)), cu(0.02f0 .* randn(Float32, 3, B)))
    xyz_tgt = M.get_tgt_xyz_from_plane_disparity(xyz_src, pose)
    @test size(xyz_tgt) == (3, W, H, Np, B)
    rgb, sigma = CUDA.rand(Float32, W, H, 3, Np, B), CUDA.rand(Float32, W, H, 1, Np, B)
    rgb_out, tacc, wts = M.plane_volume_rendering(rgb, sigma, xyz_tgt)
    @test size(rgb_out) == (W, H, 3, B) && all(Array(tacc)[:, :, :, 1, :] .== 1)
    out, depth, mask = M.render_tgt_rgb_depth(rgb, sigma, disparity, xyz_tgt, pose, K_inv, K)
    @test size(out) == (W, H, 3, B) && size(depth) == (W, H, 1, Np, B) && size(mask) == (W, H, 1, 1, B)
    g = Zygote.gradient(s -> sum(M.render_tgt_rgb_depth(rgb, s, disparity, xyz_tgt, pose, K_inv, K)[1]), sigma)[1]
    @test size(g) == size(sigma) && all(isfinite, Array(g))
end

@testset "dataset ingestion on the device" begin
    # Depth10k (src/dtk.jl:29-46): the decoder's row-major RGB raster of a triplet image, (3, 3W, H, N) in Julia's order
    W, H, N = 20, 12, 2
    strips = rand(UInt8, 3, 3W, H, N)
    x = M.MDB200.depth10k_batch(cu(strips))
    @test size(x) == (W, H, 3, 3, N)
    ref = [Float32(reinterpret(N0f8, strips[c, (l - 1) * W + w, h, n])) for w in 1:W, h in 1:H, c in 1:3, l in 1:3, n in 1:N]
    @test Array(x) == ref                                   # Float32(::N0f8), bit for bit
    xu = M.MDB200.depth10k_batch_u8(cu(strips))
    @test eltype(xu) == UInt8 && Float32.(reinterpret.(N0f8, Array(xu))) == ref
    flips = cu(UInt8[1, 0])
    xf = M.MDB200.depth10k_batch(cu(strips); flips = flips)
    @test Array(xf)[:, :, :, :, 1] == ref[end:-1:1, :, :, :, 1] && Array(xf)[:, :, :, :, 2] == ref[:, :, :, :, 2]
    # KittyDataset (src/kitty.jl:45-61): three gray frames per sample, imresize + Float32 on the device
    Wo, Ho = 62, 19
    frames = rand(UInt8, Wo, Ho, 3, N)
    xk = M.MDB200.kitti_batch(cu(frames), W, H)
    @test size(xk) == (W, H, 1, 3, N)
    img = reinterpret.(Gray{N0f8}, permutedims(frames[:, :, 2, 1], (2, 1)))              # what `load` returns: (rows, cols)
    want = permutedims(Float32.(channelview(imresize(img, (H, W)))), (2, 1))
    @test Array(xk)[:, :, 1, 2, 1] == want                  # pins the restatement of ImageTransformations.imresize
end


@testset "network-side operators" begin
    # PoseDecoder tail (src/pose_decoder.jl:29-31) and its pullback
    w, h, N = 5, 3, 4
    conv = CUDA.randn(Float32, w, h, 6, N)
    p = M.MDB200.pose_head(conv)
    ref = 1f-2 .* reshape(mean(Array(conv); dims = (1, 2)), (6, 1, N))
    @test Array(p.rvec) ≈ ref[1:3, 1, :] rtol = 1e-6
    @test Array(p.tvec) ≈ ref[4:6, :, :] rtol = 1e-6
    g = Zygote.gradient(c -> (q = M.MDB200.pose_head(c); sum(q.rvec) + 2 * sum(q.tvec)), conv)[1]
    @test Array(g)[:, :, 1:3, :] ≈ fill(1f-2 / (w * h), w, h, 3, N) && Array(g)[:, :, 4:6, :] ≈ fill(2f-2 / (w * h), w, h, 3, N)
    # DepthDecoder heads (src/depth_decoder.jl:46): σ of every level, and σ' from the saved output
    logits = [CUDA.randn(Float32, 8 * 2^k, 4 * 2^k, 1, N) .* 8 for k in 0:3]
    disp = M.MDB200.disparity_heads(logits)
    @test all(Array(disp[k]) ≈ Flux.σ.(Array(logits[k])) for k in 1:4)
    gl = Zygote.gradient(l -> sum(sum, M.MDB200.disparity_heads(l)), logits)[1]
    @test all(Array(gl[k]) ≈ (s = Flux.σ.(Array(logits[k])); s .* (1 .- s)) for k in 1:4)
    # find_static's statistic (src/dtk.jl:58-60), a whole batch at once
    x = CUDA.rand(Float32, 32, 16, 3, 3, N)
    s = M.MDB200.static_scores(M.SSIM(), x, 2; source_ids = [1, 3])
    want = [mean(Array(M.automasking_loss(M.SSIM(), x[:, :, :, :, n:n], x[:, :, :, 2, n:n]; source_ids = [1, 3]))) for n in 1:N]
    @test s ≈ want rtol = 1e-5
    # grid_sample as both call sites use it
    grid = CUDA.rand(Float32, 2, 7, 5, N) .* 2.4f0 .- 1.2f0
    inp = CUDA.rand(Float32, 32, 16, 3, N)
    @test Array(M.MDB200.grid_sample_border(inp, grid)) ≈ Array(NNlib.grid_sample(inp, grid; padding_mode = :border)) atol = 1e-6
end

end
