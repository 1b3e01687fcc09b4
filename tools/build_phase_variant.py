#!/usr/bin/env python
"""Builds libmuninn_cuda_phase.so: the shipped sources copied to gpurun_out/phase_src (scratch, not tracked) with clock64()
phase counters added to the leader warp of hnsw_beam_kernel — set-up, descent, pop, list, visited, hand-over, issue,
rows + distances, wait for the team, merge — returned in out_dist[0..10) (k = 11: entry 10 = total cycles). For
tools/phase_probe.py (MUNINN_CUDA_SO=…/libmuninn_cuda_phase.so); results: profiles/r2_bench/lone_query_phases*.jsonl.
A clock read after a barrier records its issue, not its release, so time spent waiting at a barrier lands in the phase
that first touches barrier-protected data."""
import shutil, os, subprocess
R=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); D=R+'/gpurun_out/phase_src'
shutil.rmtree(D, ignore_errors=True); os.makedirs(D+'/sqlite_muninn_b200')
shutil.copytree(R+'/sqlite_muninn_b200/csrc', D+'/sqlite_muninn_b200/csrc', ignore=shutil.ignore_patterns('*.so'))
shutil.copytree(R+'/include', D+'/include')
p=D+'/sqlite_muninn_b200/csrc/hnsw_beam.cuh'
s=open(p).read()
def rp(a,b,c=1):
    global s
    assert s.count(a)==c,(a,s.count(a)); s=s.replace(a,b)
rp('''        long long n_dist = 0, n_dist_upper = 0;
''','''        long long n_dist = 0, n_dist_upper = 0;
        long long T[10] = {0,0,0,0,0,0,0,0,0,0}; long long tc = clock64(), tq0 = tc;
#define PH(i) do { long long now_ = clock64(); T[i] += now_ - tc; tc = now_; } while (0)
''')
rp('''        if (leader && p.entry_slot >= 0 && p.n_live > 0) {
''','''        PH(0);
        if (leader && p.entry_slot >= 0 && p.n_live > 0) {
''')
rp('''            entry0 = cur;
''','''            entry0 = cur;
            PH(1);
''')
rp('''            int nf = -1;
            if (leader) {
''','''            int nf = -1;
            tc = clock64();
            if (leader) {
''')
rp('''                    const int c = (int)((unsigned)R[pos] >> 1);
                    __syncwarp();
''','''                    const int c = (int)((unsigned)R[pos] >> 1);
                    __syncwarp();
                    PH(2);
''')
rp('''                        /* lists of more than 32 entries: the second chunk's visited buckets move towards L2 while the first is probed */
''','''                        if (__any_sync(FULL, t[0] == -12345)) T[9]++;
                        PH(3);
''')
rp('''                        n_ins += nf;
                        n_dist += nf;
                        __syncwarp();
''','''                        n_ins += nf;
                        n_dist += nf;
                        __syncwarp();
                        PH(4);
''')
rp('''            if (nf < 0) break;

            if (nf > 0) {
''','''            if (nf < 0) break;
            PH(5);
            if (nf > 0) {
''')
rp('''            if (nf > 0) ring.run(p, q, Fk, qn2, lane, pol);
            team_sync<W>();
''','''            PH(6);
            if (nf > 0) ring.run(p, q, Fk, qn2, lane, pol);
            PH(7);
            team_sync<W>();
            PH(8);
''')
rp('''                stale = improved ? 0 : stale + 1; /* :428-432 */
            }
''','''                stale = improved ? 0 : stale + 1; /* :428-432 */
                PH(9);
            }
''')
rp('''                    p.out_dist[(size_t)qi * p.k + i] = d;
''','''                    p.out_dist[(size_t)qi * p.k + i] = i < 10 ? (float)T[i] : (float)(clock64() - tq0);
''')
open(p,'w').write(s)
if os.environ.get("PHASE_VISITED"):
    # optional: split the visited test-and-set further (bucket load / match.any / the rest) into out_dist[11..14), k >= 14
    p2=D+'/sqlite_muninn_b200/csrc/hnsw_search.cuh'
    t=open(p2).read()
    def rp2(a,b,c=1):
        global t
        assert t.count(a)==c,(a,t.count(a)); t=t.replace(a,b)
    rp2('__device__ __forceinline__ int visited_test_and_set(', '__device__ long long g_vis_dbg[4];\n__device__ __forceinline__ int visited_test_and_set(')
    rp2('    while (__any_sync(FULL, active)) {\n        bool claim = false;', '    while (__any_sync(FULL, active)) {\n        long long v0_ = clock64();\n        bool claim = false;')
    rp2('        const unsigned cm = __ballot_sync(FULL, claim);\n        if (claim) {\n            const unsigned peers = __match_any_sync(cm, b);',
        '        const unsigned cm = __ballot_sync(FULL, claim);\n        long long v1_ = clock64();\n        if (lane == 0) { g_vis_dbg[0] += v1_ - v0_; g_vis_dbg[3] += 1; }\n        if (claim) {\n            const unsigned peers = __match_any_sync(cm, b);\n            if (peers == 0xdeadbeefu) g_vis_dbg[2]++;\n            { long long v2_ = clock64(); if (lane == 0 || (cm & ((1u << lane) - 1)) == 0) g_vis_dbg[1] += v2_ - v1_; }')
    open(p2,'w').write(t)
    rp("p.out_dist[(size_t)qi * p.k + i] = i < 10 ? (float)T[i] : (float)(clock64() - tq0);",
       "p.out_dist[(size_t)qi * p.k + i] = i < 10 ? (float)T[i] : i == 10 ? (float)(clock64() - tq0) : (float)g_vis_dbg[(i - 11) & 3];")
    open(p,'w').write(s)
subprocess.check_call('nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -o %s/sqlite_muninn_b200/csrc/libmuninn_cuda_phase.so muninn_cuda.cu' % R, shell=True, cwd=D+'/sqlite_muninn_b200/csrc')
