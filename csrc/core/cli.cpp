// Command-line grammar (layer L1).  Same five positionals, same error text and
// same output-name rule as the reference programs:
//   Usage()          mpi/mpi_convolution.c:328-348
//   Initialization() cuda/functions.c:10-29
//   "blur_"+argv[1]  mpi_convolution.c:244-246, cuda/main.c:36-38
// Trailing --flags are extensions (SURVEY §5 "Config / flag system"); each one
// has a B200CONV_* environment mirror that the flag overrides.
#include <cstdlib>

#include "common.h"

namespace {

int parse_boundary(const char *s) {
    if (!strcmp(s, "zero")) return B2C_ZERO;
    if (!strcmp(s, "torus")) return B2C_TORUS;
    if (!strcmp(s, "cuda-compat") || !strcmp(s, "cuda_compat")) return B2C_CUDA_COMPAT;
    return -1;
}
int parse_filter(const char *s) {
    if (!strcmp(s, "gaussian")) return B2C_GAUSSIAN;
    if (!strcmp(s, "box")) return B2C_BOX;
    if (!strcmp(s, "edge")) return B2C_EDGE;
    return -1;
}
int parse_kernel(const char *s) {
    if (!strcmp(s, "auto")) return B2C_KERNEL_AUTO;
    if (!strcmp(s, "naive")) return B2C_KERNEL_NAIVE;
    if (!strcmp(s, "stream")) return B2C_KERNEL_STREAM;
    if (!strcmp(s, "tma")) return B2C_KERNEL_TMA;
    if (!strcmp(s, "resident")) return B2C_KERNEL_RESIDENT;
    return -1;
}

// Applies one "name value" option; returns false on a bad name or value.
bool apply_option(b2c_cli *c, const char *name, const char *val) {
    if (!strcmp(name, "gpus")) return (c->gpus = atoi(val)) >= 0;
    if (!strcmp(name, "boundary")) return (c->boundary = parse_boundary(val)) >= 0;
    if (!strcmp(name, "filter")) return (c->filter = parse_filter(val)) >= 0;
    if (!strcmp(name, "tb")) return (c->tb = atoi(val)) >= 0;
    if (!strcmp(name, "kernel")) return (c->kernel = parse_kernel(val)) >= 0;
    if (!strcmp(name, "check-every")) return (c->check_every = atoi(val)) >= 0;
    if (!strcmp(name, "halo-depth")) return (c->halo_depth = atoi(val)) >= 0;
    if (!strcmp(name, "checkpoint-every")) return (c->checkpoint_every = atoi(val)) >= 0;
    if (!strcmp(name, "json")) { c->json = val; return true; }
    if (!strcmp(name, "cpu")) {   // host fallback: thread count, or "all"
        c->cpu = !strcmp(val, "all") ? -1 : atoi(val);
        return c->cpu != 0 && c->cpu >= -1;
    }
    return false;
}

void apply_env(b2c_cli *c) {
    static const char *names[][2] = {
        {"B200CONV_GPUS", "gpus"},           {"B200CONV_BOUNDARY", "boundary"},
        {"B200CONV_FILTER", "filter"},       {"B200CONV_TB", "tb"},
        {"B200CONV_KERNEL", "kernel"},       {"B200CONV_CHECK_EVERY", "check-every"},
        {"B200CONV_CHECKPOINT_EVERY", "checkpoint-every"}, {"B200CONV_JSON", "json"},
        {"B200CONV_CPU", "cpu"},             {"B200CONV_HALO_DEPTH", "halo-depth"}};
    for (auto &n : names) {
        const char *v = getenv(n[0]);
        if (v && *v) apply_option(c, n[1], v);
    }
    const char *p = getenv("B200CONV_PROFILE");
    if (p && *p && strcmp(p, "0")) c->profile = 1;
}

}  // namespace

extern "C" {

int b2c_parse_cli(int argc, char **argv, b2c_cli *out) {
    if (!out) return B2C_EINVAL;
    memset(out, 0, sizeof *out);
    if (argc < 6) {
        b2c::set_error("Error Input!");
        return B2C_EINVAL;
    }
    if (!strcmp(argv[5], "grey"))
        out->channels = 1;
    else if (!strcmp(argv[5], "rgb"))
        out->channels = 3;
    else {
        b2c::set_error("Error Input!");
        return B2C_EINVAL;
    }
    out->image = argv[1];
    out->width = atoi(argv[2]);   // atoi, no range check, like the reference
    out->height = atoi(argv[3]);
    out->repetitions = atoi(argv[4]);
    apply_env(out);
    for (int i = 6; i < argc; ++i) {
        const char *a = argv[i];
        if (strncmp(a, "--", 2) != 0) {
            b2c::set_error("Error Input!");
            return B2C_EINVAL;
        }
        if (!strcmp(a, "--profile")) {
            out->profile = 1;
            continue;
        }
        if (i + 1 >= argc || !apply_option(out, a + 2, argv[i + 1])) {
            b2c::set_error("Error Input! bad option %s", a);
            return B2C_EINVAL;
        }
        ++i;
    }
    return B2C_OK;
}

int b2c_usage_text(const char *argv0, int style, char *buf, size_t buflen) {
    int n;
    if (style == 0)
        n = snprintf(buf, buflen, "\nError Input!\n%s Image_name width height repetitions [rgb/grey].\n\n", argv0);
    else
        n = snprintf(buf, buflen, "Error Input!\n%s image_name width height repetitions [rgb/grey].\n", argv0);
    return (n < 0 || (size_t)n >= buflen) ? B2C_EINVAL : B2C_OK;
}

int b2c_output_name(const char *image, char *buf, size_t buflen) {
    int n = snprintf(buf, buflen, "blur_%s", image);
    return (n < 0 || (size_t)n >= buflen) ? B2C_EINVAL : B2C_OK;
}

}  // extern "C"
