#!/usr/bin/env Rscript
# Runs the UNMODIFIED sequenza package (cran/sequenza 3.0.0) on this repository's synthetic
# inputs and writes its outputs at full precision, so that the CPU oracle and the CUDA path
# can be compared with the real reference the day a machine has R (none of the build or GPU
# boxes has: `Rscript` is absent, which is why parity is "unpinned" against R today).
#
#   Rscript tools/r_reference_dump.R --dump tests/golden/r_inputs tests/golden
#       reads   <in>/baf_segments.tsv, <in>/mufreq_positions.tsv, <in>/grid.json
#               (written by tools/make_r_inputs.py)
#       writes  <out>/r_baf_bayes_cn{7,20}_rp{0,1}.json, r_baf_model_fit.json, r_get_ci.json,
#               r_mufreq_bayes_{scalar,vector}.json, r_mufreq_model_fit.json,
#               r_session.json   (%.17g numbers)
#       tests/test_r_golden.py then checks oracle and GPU against those files.
#
#   Rscript tools/r_reference_dump.R --time-grid segments.tsv CE PL CNT_MAX CORES STEPS WARMUP
#       times baf.model.fit(mc.cores = CORES) on the comma-separated grid axes CE x PL and
#       prints the mean seconds per call on the last line (bench.py --impl reference uses
#       this when Rscript and the package are installed; kind "reference-R").
#
# Requires: R >= 3.2.0 with sequenza 3.0.0 and its imports (pbapply, copynumber, ...).
suppressMessages(library(sequenza))

num <- function(x) {
    s <- sprintf("%.17g", x)
    s[is.na(x) & !is.nan(x)] <- "\"NA\""
    s[is.nan(x)] <- "\"NaN\""
    s[is.infinite(x) & x > 0] <- "\"Inf\""
    s[is.infinite(x) & x < 0] <- "\"-Inf\""
    paste0("[", paste(s, collapse = ","), "]")
}
write.json <- function(path, fields) {
    body <- paste(sprintf("\"%s\": %s", names(fields), unlist(fields)), collapse = ",\n ")
    writeLines(paste0("{\n ", body, "\n}"), path)
}
read.grid <- function(path) {
    txt <- paste(readLines(path), collapse = " ")
    get <- function(key) {
        m <- regmatches(txt, regexpr(sprintf("\"%s\"\\s*:\\s*\\[[^]]*\\]", key), txt))
        as.numeric(strsplit(gsub("^.*\\[|\\].*$", "", m), ",")[[1]])
    }
    list(cellularity = get("cellularity"), ploidy = get("ploidy"), point = get("point"))
}

args <- commandArgs(trailingOnly = TRUE)
if (length(args) >= 1 && args[1] == "--time-grid") {
    seg <- read.delim(args[2])
    ce <- as.numeric(strsplit(args[3], ",")[[1]])
    pl <- as.numeric(strsplit(args[4], ",")[[1]])
    cn <- as.integer(args[5]); cores <- as.integer(args[6])
    steps <- as.integer(args[7]); warm <- as.integer(args[8])
    call <- function() baf.model.fit(cellularity = ce, ploidy = pl, mc.cores = cores,
        Bf = seg$Bf, depth.ratio = seg$depth_ratio, avg.depth.ratio = 1,
        sd.Bf = seg$sd_Bf, sd.ratio = seg$sd_ratio, weight.Bf = seg$weight_Bf,
        weight.ratio = seg$weight_ratio, CNt.max = cn,
        priors.table = data.frame(CN = 2, value = 2))
    for (i in seq_len(warm)) invisible(call())
    t <- system.time(for (i in seq_len(steps)) invisible(call()))[["elapsed"]]
    cat(sprintf("%.6f\n", t / steps))
    quit(status = 0)
}
if (length(args) < 3 || args[1] != "--dump") stop("usage: see the header of this file")
indir <- args[2]; outdir <- args[3]
seg <- read.delim(file.path(indir, "baf_segments.tsv"))
mut <- read.delim(file.path(indir, "mufreq_positions.tsv"))
g <- read.grid(file.path(indir, "grid.json"))
pri <- data.frame(CN = 2, value = 2)   # sequenza.fit's default priors.table (R/fit.R:6)
cores <- getOption("mc.cores", 2L)

# baf.bayes at one point, both CNt.max values and both ratio.priority settings
for (cn in c(7, 20)) for (rp in c(FALSE, TRUE)) {
    b <- baf.bayes(Bf = seg$Bf, depth.ratio = seg$depth_ratio, cellularity = g$point[1],
        ploidy = g$point[2], avg.depth.ratio = 1, sd.Bf = seg$sd_Bf, sd.ratio = seg$sd_ratio,
        weight.Bf = seg$weight_Bf, weight.ratio = seg$weight_ratio, CNt.max = cn,
        priors.table = pri, ratio.priority = rp)
    write.json(file.path(outdir, sprintf("r_baf_bayes_cn%d_rp%d.json", cn, as.integer(rp))),
        list(CNt = num(b[, "CNt"]), A = num(b[, "A"]), B = num(b[, "B"]), LPP = num(b[, "LPP"])))
}
# baf.model.fit on the grid (lpp is [ploidy x cellularity], written column-major)
fit <- baf.model.fit(cellularity = g$cellularity, ploidy = g$ploidy, mc.cores = cores,
    Bf = seg$Bf, depth.ratio = seg$depth_ratio, avg.depth.ratio = 1, sd.Bf = seg$sd_Bf,
    sd.ratio = seg$sd_ratio, weight.Bf = seg$weight_Bf, weight.ratio = seg$weight_ratio,
    CNt.max = 7, priors.table = pri)
write.json(file.path(outdir, "r_baf_model_fit.json"),
    list(ploidy = num(fit$ploidy), cellularity = num(fit$cellularity),
         lpp = num(as.vector(fit$lpp))))
ci <- get.ci(fit)
write.json(file.path(outdir, "r_get_ci.json"),
    list(max_ploidy = num(ci$max.ploidy), max_cellularity = num(ci$max.cellularity),
         confint_ploidy = num(ci$confint.ploidy),
         confint_cellularity = num(ci$confint.cellularity)))
# mufreq.bayes (scalar weights as sequenza.results calls it, and per-position weights)
m <- mufreq.bayes(mufreq = mut$mufreq, depth.ratio = mut$depth_ratio,
    cellularity = g$point[1], ploidy = g$point[2], avg.depth.ratio = 1, CNt.max = 20)
write.json(file.path(outdir, "r_mufreq_bayes_scalar.json"),
    list(CNn = num(m[, "CNn"]), CNt = num(m[, "CNt"]), Mt = num(m[, "Mt"]), LPP = num(m[, "LPP"])))
m <- mufreq.bayes(mufreq = mut$mufreq, depth.ratio = mut$depth_ratio,
    cellularity = g$point[1], ploidy = g$point[2], avg.depth.ratio = 1,
    weight.mufreq = mut$weight, weight.ratio = 2 * mut$weight, CNt.max = 7)
write.json(file.path(outdir, "r_mufreq_bayes_vector.json"),
    list(CNn = num(m[, "CNn"]), CNt = num(m[, "CNt"]), Mt = num(m[, "Mt"]), LPP = num(m[, "LPP"])))
mf <- mufreq.model.fit(cellularity = g$cellularity, ploidy = g$ploidy, mc.cores = cores,
    mufreq = mut$mufreq, depth.ratio = mut$depth_ratio, avg.depth.ratio = 1,
    weight.mufreq = mut$weight, weight.ratio = 2 * mut$weight, priors.table = pri)
write.json(file.path(outdir, "r_mufreq_model_fit.json"),
    list(ploidy = num(mf$ploidy), cellularity = num(mf$cellularity),
         lpp = num(as.vector(mf$lpp))))
writeLines(sprintf("{\"R\": \"%s\", \"sequenza\": \"%s\"}", R.version.string,
    as.character(packageVersion("sequenza"))), file.path(outdir, "r_session.json"))
