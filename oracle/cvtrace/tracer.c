// tracer.c — TEST-INFRASTRUCTURE tool (never part of the product; nothing under forest_slam_b200/ uses it).
// Single-steps a child python process between two SIGUSR1 markers and records, for every instruction executed
// inside cv2.abi3.so, the instruction pointer, the general-purpose registers and xmm0-15.  Used to read the OPERATION
// ORDER of OpenCV's five-point solver (calib3d/five-point.cpp, Apache-2.0; its source is not on this machine) out of
// the cv2 4.13.0 binary the oracle is pinned to; see DESIGN.md section 2 and oracle/cvtrace/README.md.
//   usage: tracer out.bin python script.py [args]
#define _GNU_SOURCE
#include <sys/ptrace.h>
#include <sys/wait.h>
#include <sys/user.h>
#include <unistd.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <signal.h>
#include <stdint.h>

typedef struct { uint64_t rip; uint64_t gpr[16]; uint8_t xmm[16][16]; } Rec;

static int find_cv2(pid_t pid, uint64_t* lo, uint64_t* hi, FILE* mapsout) {
    char path[64], line[1024];
    snprintf(path, sizeof path, "/proc/%d/maps", pid);
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    *lo = ~0ull; *hi = 0;
    while (fgets(line, sizeof line, f)) {
        fputs(line, mapsout);
        if (strstr(line, "cv2.abi3.so") && strstr(line, "r-xp")) {
            uint64_t a, b;
            sscanf(line, "%lx-%lx", &a, &b);
            if (a < *lo) *lo = a;
            if (b > *hi) *hi = b;
        }
    }
    fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    pid_t p = fork();
    if (!p) {
        ptrace(PTRACE_TRACEME, 0, 0, 0);
        execvp(argv[2], argv + 2);
        perror("exec");
        _exit(127);
    }
    int st;
    waitpid(p, &st, 0);  // exec stop
    FILE* out = fopen(argv[1], "wb");
    char mp[512];
    snprintf(mp, sizeof mp, "%s.maps", argv[1]);
    int tracing = 0;
    uint64_t lo = 0, hi = 0;
    long total = 0, kept = 0;
    int sig = 0;
    for (;;) {
        if (ptrace(tracing ? PTRACE_SINGLESTEP : PTRACE_CONT, p, 0, sig) < 0) { perror("ptrace"); break; }
        sig = 0;
        waitpid(p, &st, 0);
        if (WIFEXITED(st) || WIFSIGNALED(st)) break;
        int s = WSTOPSIG(st);
        if (s == SIGUSR1) {
            if (!tracing) {
                FILE* mo = fopen(mp, "w");
                find_cv2(p, &lo, &hi, mo);
                fclose(mo);
                fprintf(stderr, "tracing on: cv2 text %lx-%lx\n", lo, hi);
                tracing = 1;
            } else {
                fprintf(stderr, "tracing off: %ld steps, %ld kept\n", total, kept);
                tracing = 0;
            }
            continue;
        }
        if (s != SIGTRAP) { sig = s; continue; }
        if (!tracing) continue;
        total++;
        struct user_regs_struct r;
        ptrace(PTRACE_GETREGS, p, 0, &r);
        if (r.rip < lo || r.rip >= hi) continue;
        struct user_fpregs_struct fr;
        ptrace(PTRACE_GETFPREGS, p, 0, &fr);
        Rec rec;
        rec.rip = r.rip;
        uint64_t g[16] = {r.rax, r.rcx, r.rdx, r.rbx, r.rsp, r.rbp, r.rsi, r.rdi,
                          r.r8, r.r9, r.r10, r.r11, r.r12, r.r13, r.r14, r.r15};
        memcpy(rec.gpr, g, sizeof g);
        memcpy(rec.xmm, fr.xmm_space, 256);
        fwrite(&rec, sizeof rec, 1, out);
        kept++;
    }
    fclose(out);
    return 0;
}
