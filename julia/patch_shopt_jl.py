#!/usr/bin/env python3
"""patch_shopt_jl.py PATH/TO/shopt/shopt.jl [-o OUT]  --  the reference-side change that binds libshopt_b200.so.

Replaces the hot loops of the reference's top-level script by one call each into julia/shopt_b200.jl (module ShoptB200,
a ccall shim over include/shopt_b200.h) and adds the include.  The replaced line ranges (1-based, inclusive, in the
UNMODIFIED file whose sha256 is checked below) are:

    79        include("reader.jl")                       + include of shopt_b200.jl
    131-164   model analytic fit loop                    -> ShoptB200.fit_analytic_batch   (cost / g!)
    194-224   mode == "chisq" pixel-grid loop            -> ShoptB200.chisq_mode
    281-314   mode == "PCA" pixel-grid loop              -> ShoptB200.pca_batch
    316-348   mode == "smooth" pixel-grid loop           -> ShoptB200.smooth_mode
    359-397   learned analytic fit loop                  -> ShoptB200.fit_analytic_batch   (costD / gD!)
    534-577   polyfit rounds                             -> ShoptB200.polyfit

Everything else (cataloging, the splices, the s/g1/g2 field fits, plots, writeFitsData) is untouched, so the script keeps
its globals, its printed log and its shopt.yml semantics.  The patch is expressed as line ranges + new text (no line of
the reference is reproduced here); it refuses to run on a file with a different checksum.  Copy julia/shopt_b200.jl
next to the patched shopt.jl (or set LIBSHOPT_B200 to the library's path).

Julia is absent from the build image: the patch is applied to a scratch copy by tests/test_julia_shim.py (ranges land on
the loops they name, blocks balance their begin/end), not executed.
"""
import hashlib
import sys

SHA256 = "cc1cea501bbdd2cdd15fc7813422b37f66fe6d87c424837d17ae5d14a5088b20"
N_LINES = 820

BLOCKS = {
    (79, 79): '''  include("reader.jl")
  include(joinpath(@__DIR__, "shopt_b200.jl"))   # libshopt_b200.so binding: the per-star fitting stage on B200''',

    (131, 164): '''@time begin
  # shopt_b200: one batched LBFGS fit of cost / g! over the catalogue (x0 = the reference's rand(3) per star)
  s_b, g1_b, g2_b, status_model, _ = ShoptB200.fit_analytic_batch(starCatalog[1:itr], AnalyticStampSize,
                                                                  minAnalyticGradientModel; x0 = rand(3, itr))
  for i in 1:itr
    s_model[i], g1_model[i], g2_model[i] = s_b[i], g1_b[i], g2_b[i]
    status_model[i] == ShoptB200.FIT_THREW && println("Star $i failed")      # the catch branch: s = g1 = g2 = 0
  end
end''',

    (194, 224): '''if mode == "chisq"
  println("━  χ2  Mode...\\n")
  @time begin
    # shopt_b200: the minimiser of chisq_cost written down (chisq.jl:1-11); rand(r, c, itr) = the reference's starts
    append!(pixelGridFits, ShoptB200.chisq_mode(starCatalog[1:itr], errVignets[1:itr], unity_sum; x0 = rand(r, c, itr)))
  end
end''',

    (281, 314): '''if mode == "PCA"
  println("━ PCA Mode...")
  @time begin
    # shopt_b200: nanToZero -> pca_image(., PCAterms) -> smooth(., lanczos) (or ./ sum when unity_sum), one call
    pgf_b, failed_b, _ = ShoptB200.pca_batch(starCatalog, PCAterms, lanczos, unity_sum)
    append!(pixelGridFits, pgf_b)
    for i in failed_b
      println("Star $i failed")
      push!(failedStars, i)
    end
  end
end''',

    (316, 348): '''if mode == "smooth"
    println("━ Smooth Mode...")
    @time begin
        # shopt_b200: nanToZero -> smooth(., lanczos) (or ./ sum when unity_sum) for every star in one call
        append!(pixelGridFits, ShoptB200.smooth_mode(starCatalog, lanczos, unity_sum))
    end
end''',

    (359, 397): '''@time begin
  # shopt_b200: one batched LBFGS fit of costD / gD! over the pixel-grid fits
  s_b, g1_b, g2_b, status_learned, _ = ShoptB200.fit_analytic_batch(pixelGridFits, AnalyticStampSize,
                                                                    minAnalyticGradientLearned;
                                                                    x0 = rand(3, length(starCatalog)))
  for i in 1:length(starCatalog)
    s_data[i], g1_data[i], g2_data[i] = s_b[i], g1_b[i], g2_b[i]
    if status_learned[i] == ShoptB200.FIT_THREW            # the catch branch: zeros, NOT pushed to failedStars
      println("Star $i failed")
    elseif s_data[i] < sLowerBound || s_data[i] > sUpperBound
      push!(failedStars, i)
    end
  end
end''',

    (534, 577): '''@time begin
  # shopt_b200: the polyfit rounds (polynomial_optimizer per pixel, per-star MSE, worst_10_percent, the drop loop)
  global training_stars, training_u_coords, training_v_coords, PolynomialMatrix
  println("$(length(training_stars)) training stars")
  PolynomialMatrix, kept_training, n_per_round = ShoptB200.polyfit(pixelGridFits, u_coordinates, v_coordinates,
                                                                  Vector{Int}(training_indices), degree,
                                                                  iterationsPolyfit;
                                                                  median_constraint = median_constraint,
                                                                  outlier_sigma_clipping = outlier_sigma_clipping)
  for loop in 1:iterationsPolyfit
    println("━ Iteration: $loop")
    loop != iterationsPolyfit && println("$(n_per_round[loop + 1]) training stars")
  end
  training_stars = pixelGridFits[kept_training]
  training_u_coords = u_coordinates[kept_training]
  training_v_coords = v_coordinates[kept_training]
end''',
}


def apply(text: str) -> str:
    if hashlib.sha256(text.encode()).hexdigest() != SHA256:
        raise SystemExit("patch_shopt_jl: this is not the shopt/shopt.jl the line ranges were taken from (sha256 differs)")
    lines = text.split("\n")
    for (a, b) in sorted(BLOCKS, reverse=True):           # bottom-up: earlier ranges keep their numbering
        lines[a - 1:b] = BLOCKS[(a, b)].split("\n")
    return "\n".join(lines)


def main(argv):
    if len(argv) < 2:
        raise SystemExit(__doc__)
    out = argv[argv.index("-o") + 1] if "-o" in argv else argv[1]
    with open(argv[1], encoding="utf-8") as f:
        new = apply(f.read())
    with open(out, "w", encoding="utf-8") as f:
        f.write(new)
    print("patched %s -> %s (%d loops replaced)" % (argv[1], out, len(BLOCKS) - 1))


if __name__ == "__main__":
    main(sys.argv)
