# run_reference_on_dump.jl -- NOT executed on the build or GPU machines (no Julia there).
#
# Pins the parity this repository cannot pin itself: feeds the UNMODIFIED reference (Monodepth2.jl) the
# bytes written by `python -m mdb200.synth <config> <dir>` and writes the reference's own loss and
# gradients next to them, for `tools/compare_reference_dump.py` to hold against the oracle / the GPU.
#
#   julia --project=<Monodepth2.jl checkout> run_reference_on_dump.jl <dir>
#
# `train_loss` takes a model; the dump holds the model's OUTPUTS (disparities, poses), so the closure
# below is `train_loss`'s body from src/training.jl:27-77 applied to them (SURVEY F6).
using JSON, Zygote, Flux, Statistics
using Monodepth
using Monodepth: Pose, composeT, Backproject, Project, SSIM, disparity_to_depth, prediction_loss, smooth_loss

dir = ARGS[1]
meta = JSON.parsefile(joinpath(dir, "inputs.json"))
rd(name) = reshape(reinterpret(Float32, read(joinpath(dir, name * ".f32"))), Tuple(meta["julia_shapes"][name])) |> collect

x = rd("x")
W, H, C, L, N = size(x)
Ls = length(meta["scales"])
disparities = [rd("disp$(k - 1)") for k in 1:Ls]
source_ids, target_id = Int.(meta["source_ids"]), Int(meta["target_id"])
poses = [Pose(rd("rvec$(s - 1)"), rd("tvec$(s - 1)")) for s in 1:length(source_ids)]
K, invK = rd("K"), rd("invK")
ssim, backproject, project = SSIM(), Backproject(; width=W, height=H), Project(; width=W, height=H)
target = x[:, :, :, target_id, :]

function loss_of(disparities, poses)                      # src/training.jl:27-77 on given network outputs
    Ps = map(p -> composeT(p[1].rvec, p[1].tvec, p[2]), zip(poses, source_ids .< target_id))
    loss = 0f0
    for (disparity, scale) in zip(disparities, Float32.(meta["scales"]))
        d = size(disparity)[1:2] == (W, H) ? disparity : upsample_bilinear(disparity; size=(W, H))
        depth = reshape(disparity_to_depth(d, Float32(meta["min_depth"]), Float32(meta["max_depth"])), (1, W * H, N))
        coords = backproject(depth, invK)
        warped = map(zip(Ps, source_ids)) do (P, sid)
            uvs = reshape(project(coords, K, P...), (2, W, H, N))
            grid_sample(x[:, :, :, sid, :], uvs; padding_mode=:border)
        end
        warp_loss = prediction_loss(ssim, warped, target)
        nd = d ./ (mean(d; dims=(1, 2)) .+ 1f-7)
        loss += mean(warp_loss) + smooth_loss(reshape(nd, (W, H, N)), target) * Float32(meta["disparity_smoothness"]) * scale
    end
    loss / Ls
end

val, back = Zygote.pullback(loss_of, disparities, poses)
gd, gp = back(1f0)
write(joinpath(dir, "ref_loss.f32"), Float32[val])
for k in 1:Ls
    write(joinpath(dir, "ref_d_disp$(k - 1).f32"), Float32.(gd[k]))
end
for s in 1:length(poses)
    write(joinpath(dir, "ref_d_rvec$(s - 1).f32"), Float32.(gp[s].rvec))
    write(joinpath(dir, "ref_d_tvec$(s - 1).f32"), Float32.(gp[s].tvec))
end
println("reference loss = ", val)
