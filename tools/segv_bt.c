// LD_PRELOAD helper (dev only): print a native backtrace on SIGSEGV.
#define _GNU_SOURCE
#include <execinfo.h>
#include <signal.h>
#include <stdio.h>
#include <unistd.h>
static void on_segv(int sig, siginfo_t* si, void* u) {
    void* bt[64];
    int n = backtrace(bt, 64);
    fprintf(stderr, "\n[segv_bt] signal %d at address %p\n", sig, si->si_addr);
    backtrace_symbols_fd(bt, n, 2);
    _exit(139);
}
__attribute__((constructor)) static void init(void) {
    struct sigaction sa = {0};
    sa.sa_sigaction = on_segv;
    sa.sa_flags = SA_SIGINFO | SA_ONSTACK;
    sigaction(SIGSEGV, &sa, 0);
}
