/* sprof.c -- a 60-line sampling profiler for boxes without perf/gdb (measurement helper, not product):
 *   gcc -O1 -shared -fPIC -o sprof.so sprof.c -ldl;  LD_PRELOAD=./sprof.so SPROF_OUT=out.txt <server> ...
 * SIGPROF every 2 ms of process CPU time; per sample the thread's name, the module and offset of the PC and of the
 * word on top of the stack (the caller, for leaf routines); written when the process exits (SIGTERM -> exit).
 * Offsets are symbolised afterwards with addr2line against the same binary. */
#define _GNU_SOURCE
#include <signal.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <dlfcn.h>
#include <sys/time.h>
#include <ucontext.h>
#include <unistd.h>
#define MAXS (1<<22)
#include <sys/syscall.h>
static void *pcs[MAXS]; static void *ras[MAXS]; static int tids[MAXS]; static volatile long n;
static void onprof(int s, siginfo_t *si, void *u) {
	ucontext_t *uc = (ucontext_t*)u;
	long i = __sync_fetch_and_add(&n, 1);
	if(i < MAXS) { pcs[i] = (void*)uc->uc_mcontext.gregs[REG_RIP]; ras[i] = *(void**)uc->uc_mcontext.gregs[REG_RSP]; tids[i] = (int)syscall(SYS_gettid); }
}
static void dump(void) {
	FILE *f = fopen(getenv("SPROF_OUT") ? getenv("SPROF_OUT") : "/tmp/sprof.txt", "w");
	long m = n < MAXS ? n : MAXS;
	static char names[1<<22][8]; /* tid -> first 8 chars of comm */
	for(long i = 0; i < m; i++) {
		int t = tids[i] & ((1<<22)-1);
		if(!names[t][0]) { char p[64]; snprintf(p, sizeof p, "/proc/self/task/%d/comm", tids[i]); FILE *c = fopen(p, "r"); if(c) { if(fgets(names[t], 8, c)) { char *q = strchr(names[t], '\n'); if(q) *q = 0; } fclose(c); } else strcpy(names[t], "gone"); }
	}
	for(long i = 0; i < m; i++) {
		Dl_info d, e; const char *mod = "?", *sym = "?", *sym2 = "?", *mod2 = "?"; char a1[64], a2[64];
		if(dladdr(pcs[i], &d)) { if(d.dli_fname) mod = d.dli_fname; if(d.dli_sname) sym = d.dli_sname; else { snprintf(a1, sizeof a1, "+%lx", (unsigned long)((char*)pcs[i] - (char*)d.dli_fbase)); sym = a1; } }
		if(dladdr(ras[i], &e)) { if(e.dli_fname) mod2 = strrchr(e.dli_fname, '/') ? strrchr(e.dli_fname, '/') + 1 : e.dli_fname; snprintf(a2, sizeof a2, "%s+%lx", mod2, (unsigned long)((char*)ras[i] - (char*)e.dli_fbase)); sym2 = a2; }
		const char *b = strrchr(mod, '/');
		fprintf(f, "%s\t%s\t%s\t%s\n", names[tids[i] & ((1<<22)-1)], b ? b + 1 : mod, sym, sym2);
	}
	fclose(f);
}
static void onterm(int s) { exit(0); }
__attribute__((constructor)) static void init(void) {
	struct sigaction sa; memset(&sa, 0, sizeof sa);
	sa.sa_sigaction = onprof; sa.sa_flags = SA_SIGINFO | SA_RESTART;
	sigaction(SIGPROF, &sa, NULL);
	struct itimerval it = {{0, 2000}, {0, 2000}};
	setitimer(ITIMER_PROF, &it, NULL);
	signal(SIGTERM, onterm);
	atexit(dump);
}
