/*
 * driver/main.c -- C host driver with the reference's command line and output contract.
 *
 *   parallel [param_file] [block_size]
 *
 * Mirrors src/parallel/main.cu:954-1129 / src/serial/main.c:31-158 as a process:
 *   - parameter file: 15 positional lines (io.c:12-59), default "parameters.txt" (main.c:61-66);
 *   - block_size (main.cu:987-1000) is accepted for script compatibility and ignored: tile shapes
 *     are chosen by the library.  Unlike the reference it is optional (run.sh:136 and the notebook
 *     call the binary without it);
 *   - stdout: "U-CENTER: %.6f\nV-CENTER: %.6f\n" (main.c:148-149);
 *   - stderr: accumulated SOR seconds "%.6f", no newline, last thing written (main.c:153);
 *   - unknown problem type: message + EXIT_FAILURE (main.c:105-108);
 *   - field dumps every n_print-th step to out/<n> are DISABLED in both reference mains
 *     (main.c:138-143 is commented out); set NSB_OUTPUT=1 to enable them here (single GPU and NSB_GPUS alike).
 *     The dump is overlapped with the following time steps (nsb_output_async: device snapshot, side-stream copy,
 *     host formatting threads); NSB_OUTPUT=sync writes it before the loop goes on, like the reference would;
 *   - NSB_GPUS=n (n > 1) cuts the domain into n y-slabs, one per GPU of the box (new: the reference
 *     is single-GPU).
 *
 * Plain C99, no CUDA headers: everything below the C ABI of include/nsb200.h is in libnsb200.so.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nsb200.h"

int main(int argc, char* argv[])
{
    const char* param_file = "parameters.txt";
    nsb_params P;
    nsb_ctx* ctx = NULL;
    double t = 0.0;
    int n = 0, n_out = 0;
    const char* dump = getenv("NSB_OUTPUT");
    const char* verbose = getenv("NSB_VERBOSE");
    const char* gpus_env = getenv("NSB_GPUS");
    nsb_group* grp = NULL;

    if (argc > 1) param_file = argv[1];
    if (argc > 2) {
        int bs = atoi(argv[2]);
        if (bs > 0 && bs <= 32) fprintf(stderr, "CUDA: Using block size %d\n", bs);   /* main.cu:999 */
    }
    if (nsb_read_params(param_file, &P) != NSB_OK) {
        perror("Error while opening the file.\n");                                   /* io.c:18-22 */
        exit(EXIT_FAILURE);
    }
    if (P.problem != 1 && P.problem != 2) {
        printf("Unknown probem type (see parameters.txt).\n");                       /* main.c:106 */
        exit(EXIT_FAILURE);
    }
    if (gpus_env && atoi(gpus_env) > 1) {                                            /* y-slabs over NSB_GPUS devices */
        if (nsb_group_create(&P, NULL, atoi(gpus_env), NULL, &grp) != NSB_OK) {
            fprintf(stderr, "CUDA error: %s\n", nsb_group_last_error(NULL));
            exit(EXIT_FAILURE);
        }
    } else if (nsb_create(&P, NULL, &ctx) != NSB_OK) {
        fprintf(stderr, "CUDA error: %s\n", nsb_last_error(NULL));                   /* main.cu:36-43 */
        exit(EXIT_FAILURE);
    }
    while (t < P.T) {                                                                /* main.c:86 */
        nsb_step_info info;
        int rc = grp ? nsb_group_step(grp, t, &info) : nsb_step(ctx, t, &info);
        if (rc != NSB_OK) {
            fprintf(stderr, "CUDA error: %s\n", grp ? nsb_group_last_error(grp) : nsb_last_error(ctx));
            exit(EXIT_FAILURE);
        }
        if (verbose)
            fprintf(stderr, "step %d t %.10g dt %.10g umax %.6g vmax %.6g gamma %.6g normp %.6g its %d res %.6g\n",
                    n, info.t, info.dt, info.u_max, info.v_max, info.gamma, info.norm_p, info.sor_iters, info.res);
        if (dump && (dump[0] == '1' || dump[0] == 's') && P.n_print > 0 && n % P.n_print == 0) { /* main.c:138-143 */
            char prefix[64];
            const int sync = dump[0] == 's';
            snprintf(prefix, sizeof prefix, "out/%d", n_out);
            if (grp) rc = sync ? nsb_group_output(grp, t, prefix) : nsb_group_output_async(grp, t, prefix);
            else rc = sync ? nsb_output(ctx, t, prefix) : nsb_output_async(ctx, t, prefix);
            if (rc != NSB_OK) {
                perror(grp ? nsb_group_last_error(grp) : nsb_last_error(ctx));       /* io.c:83-87 */
                exit(EXIT_FAILURE);
            }
            printf("Output created!\n");                                             /* io.c:118 */
            n_out++;
        }
        t += info.dt;                                                                /* main.c:145 */
        n++;
    }
    {
        double uc = 0.0, vc = 0.0;
        if (grp) nsb_group_get_center(grp, &uc, &vc);
        else nsb_get_center(ctx, &uc, &vc);
        printf("U-CENTER: %.6f\n", uc);                                              /* main.c:148 */
        printf("V-CENTER: %.6f\n", vc);                                              /* main.c:149 */
    }
    if ((grp ? nsb_group_output_wait(grp) : nsb_output_wait(ctx)) != NSB_OK) {       /* the last dump, if still being written */
        perror(grp ? nsb_group_last_error(grp) : nsb_last_error(ctx));
        exit(EXIT_FAILURE);
    }
    fflush(stdout);
    fprintf(stderr, "%.6f", grp ? nsb_group_sor_seconds(grp) : nsb_sor_seconds(ctx)); /* main.c:153 */
    if (grp) nsb_group_destroy(grp);
    else nsb_destroy(ctx);
    return 0;
}
