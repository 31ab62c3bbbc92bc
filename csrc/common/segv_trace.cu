// Debug aid: with DMPI_TRACE set, print a native backtrace on SIGSEGV (pytest's faulthandler only shows Python frames).
#include <execinfo.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

static void dmpi_segv_handler(int sig, siginfo_t* info, void*) {
  void* frames[64];
  const int n = backtrace(frames, 64);
  fprintf(stderr, "[dmpi] signal %d at address %p, native backtrace:\n", sig, info ? info->si_addr : nullptr);
  backtrace_symbols_fd(frames, n, 2);
  _exit(139);
}

__attribute__((constructor)) static void dmpi_install_segv_trace() {
  if (!getenv("DMPI_TRACE")) return;
  struct sigaction sa;
  sa.sa_sigaction = dmpi_segv_handler;
  sigemptyset(&sa.sa_mask);
  sa.sa_flags = SA_SIGINFO;
  sigaction(SIGSEGV, &sa, nullptr);
}
