# Replacement bodies for the four functions of R/bayes.R that sit on the likelihood
# path (the CNt.max / n.gpus pass-through of sequenza.fit is a patch: see the end of this
# file).  Signatures, defaults and return values are the reference's; only the work
# between the argument recycling and the result assembly moves to the GPU.
# UNTESTED HERE (no R in the build image).
#
# NAMESPACE: add   useDynLib(sequenza, .registration = TRUE)
# DESCRIPTION: set NeedsCompilation: yes

# The reference's post-processing of result$LPP (R/bayes.R:117-123 and :187-193), kept as
# it is there: mean per (ploidy, cellularity) cell, axis values back from the dimnames,
# log-sum-exp normalisation.
.sqz.finish <- function(result) {
    cell.mean <- tapply(result$LPP, list(result$ploidy, result$cellularity), mean)
    top <- max(result$LPP, na.rm = TRUE)
    log.total <- log(sum(exp(result$LPP - top))) + top
    list(ploidy = as.numeric(rownames(cell.mean)),
        cellularity = as.numeric(colnames(cell.mean)),
        lpp = exp(cell.mean - log.total))
}

.sqz.priors <- function(priors.table) {
    list(CN = as.double(priors.table$CN), value = as.double(priors.table$value))
}

# R/bayes.R:38-101
baf.bayes <- function(Bf, depth.ratio, cellularity, ploidy, avg.depth.ratio,
    sd.Bf = 0.1, sd.ratio = 0.5, weight.Bf = 1, weight.ratio = 1, CNt.min = 0,
    CNt.max = 7, CNn = 2, priors.table = data.frame(CN = CNt.min:CNt.max,
    value = 1), ratio.priority = FALSE) {
    # scalar recycling exactly as the reference's data.frame() call (R/bayes.R:43-45)
    baf.tab <- data.frame(Bf = Bf, ratio = depth.ratio,
        sd.Bf = sd.Bf, sd.ratio = sd.ratio, weight.Bf = weight.Bf,
        weight.ratio = weight.ratio)
    p <- .sqz.priors(priors.table)
    bafs.L <- .Call(sqzr_baf_point, as.double(baf.tab$Bf), as.double(baf.tab$ratio),
        as.double(baf.tab$sd.Bf), as.double(baf.tab$sd.ratio),
        as.double(baf.tab$weight.Bf), as.double(baf.tab$weight.ratio),
        as.double(cellularity), as.double(ploidy), as.double(avg.depth.ratio),
        as.integer(CNt.min), as.integer(CNt.max), as.integer(CNn), p$CN, p$value,
        as.logical(ratio.priority))
    colnames(bafs.L) <- c("CNt", "A", "B", "LPP")   # R/bayes.R:99
    bafs.L
}

# R/bayes.R:103-124: only lines 109-115 change
baf.model.fit <- function(cellularity = seq(0.3, 1, by = 0.01),
    ploidy = seq(1, 7, by = 0.1), mc.cores = getOption("mc.cores", 2L),
    n.gpus = getOption("sequenza.gpus", 1L), Bf, depth.ratio, avg.depth.ratio,
    sd.Bf = 0.1, sd.ratio = 0.5, weight.Bf = 1, weight.ratio = 1, CNt.min = 0,
    CNt.max = 7, CNn = 2, priors.table = data.frame(CN = CNt.min:CNt.max,
    value = 1), ratio.priority = FALSE) {

    result <- expand.grid(ploidy = ploidy, cellularity = cellularity,
        KEEP.OUT.ATTRS = FALSE)
    baf.tab <- data.frame(Bf = Bf, ratio = depth.ratio,
        sd.Bf = sd.Bf, sd.ratio = sd.ratio, weight.Bf = weight.Bf,
        weight.ratio = weight.ratio)
    p <- .sqz.priors(priors.table)
    # mc.cores is kept for compatibility and ignored: there is no fork pool
    result$LPP <- .Call(sqzr_baf_grid, as.double(baf.tab$Bf), as.double(baf.tab$ratio),
        as.double(baf.tab$sd.Bf), as.double(baf.tab$sd.ratio),
        as.double(baf.tab$weight.Bf), as.double(baf.tab$weight.ratio),
        as.double(ploidy), as.double(cellularity), as.double(avg.depth.ratio),
        as.integer(CNt.min), as.integer(CNt.max), as.integer(CNn), p$CN, p$value,
        as.logical(ratio.priority), as.integer(n.gpus), 0L)
    .sqz.finish(result)
}

# R/bayes.R:126-173
mufreq.bayes <- function(mufreq, depth.ratio, cellularity,
    ploidy, avg.depth.ratio, weight.mufreq = 100, weight.ratio = 100,
    CNt.min = 1, CNt.max = 7, CNn = 2,
    priors.table = data.frame(CN = CNt.min:CNt.max, value = 1)) {
    p <- .sqz.priors(priors.table)
    if (length(weight.mufreq) == 1 && length(weight.ratio) == 1 &&
        weight.mufreq == round(weight.mufreq) && weight.ratio == round(weight.ratio) &&
        weight.mufreq >= 0 && weight.ratio >= 0 && weight.mufreq <= 4096 &&
        weight.ratio <= 4096 && (weight.mufreq + 1) * (weight.ratio + 1) <= 4e6) {
        # scalar weights (the defaults; R/results.R:153-157): table-driven kernel
        mufreq.tab <- data.frame(F = mufreq, ratio = depth.ratio)
        types.L <- .Call(sqzr_mufreq_point_scalar, as.double(mufreq.tab$F),
            as.double(mufreq.tab$ratio), as.double(weight.mufreq),
            as.double(weight.ratio), as.double(cellularity), as.double(ploidy),
            as.double(avg.depth.ratio), as.integer(CNt.min), as.integer(CNt.max),
            as.integer(CNn), p$CN, p$value)
        colnames(types.L) <- c("CNn", "CNt", "Mt", "LPP")
        return(types.L)
    }
    mufreq.tab <- data.frame(F = mufreq, ratio = depth.ratio,
        weight.mufreq = weight.mufreq, weight.ratio = weight.ratio)
    types.L <- .Call(sqzr_mufreq_point, as.double(mufreq.tab$F),
        as.double(mufreq.tab$ratio), as.double(mufreq.tab$weight.mufreq),
        as.double(mufreq.tab$weight.ratio), as.double(cellularity), as.double(ploidy),
        as.double(avg.depth.ratio), as.integer(CNt.min), as.integer(CNt.max),
        as.integer(CNn), p$CN, p$value)
    colnames(types.L) <- c("CNn", "CNt", "Mt", "LPP")   # R/bayes.R:171
    types.L
}

# R/bayes.R:175-194: only lines 179-185 change
mufreq.model.fit <- function(cellularity = seq(0.3, 1, by = 0.01),
    ploidy = seq(1, 7, by = 0.1),  mc.cores = getOption("mc.cores", 2L),
    n.gpus = getOption("sequenza.gpus", 1L), mufreq, depth.ratio, avg.depth.ratio,
    weight.mufreq = 100, weight.ratio = 100, CNt.min = 1, CNt.max = 7, CNn = 2,
    priors.table = data.frame(CN = CNt.min:CNt.max, value = 1)) {
    result <- expand.grid(ploidy = ploidy, cellularity = cellularity,
        KEEP.OUT.ATTRS = FALSE)
    mufreq.tab <- data.frame(F = mufreq, ratio = depth.ratio,
        weight.mufreq = weight.mufreq, weight.ratio = weight.ratio)
    p <- .sqz.priors(priors.table)
    result$LPP <- .Call(sqzr_mufreq_grid, as.double(mufreq.tab$F),
        as.double(mufreq.tab$ratio), as.double(mufreq.tab$weight.mufreq),
        as.double(mufreq.tab$weight.ratio), as.double(ploidy), as.double(cellularity),
        as.double(avg.depth.ratio), as.integer(CNt.min), as.integer(CNt.max),
        as.integer(CNn), p$CN, p$value, as.integer(n.gpus))
    .sqz.finish(result)
}

# sequenza.fit (R/fit.R:1-69): two optional arguments at the end of its signature,
#     CNt.max = 7, n.gpus = getOption("sequenza.gpus", 1L)
# forwarded to baf.model.fit (R/fit.R:45-50) and mufreq.model.fit (R/fit.R:61-65).  The edit
# is r-package/patches/sequenza-3.0.0-gpu.patch (apply with `patch -p1` in the package
# root; it also adds useDynLib to NAMESPACE and sets NeedsCompilation in DESCRIPTION).
# The reference cannot reach CNt.max = 20 from sequenza.fit; the defaults keep its result.
# get.ci / alternative.cp.solutions (R/solutions.R) and sequenza.results (R/results.R)
# need no change: they consume list(ploidy, cellularity, lpp) and the N x 4 matrices.
