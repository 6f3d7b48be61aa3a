import subprocess, json, sys, os, csv, io
sys.path.insert(0,'.')
import bench
sha=bench.kernel_source_sha()
def summ(rep):
    return subprocess.run(['python','tools/ncu_summary.py',rep],capture_output=True,text=True).stdout
HEAD="`tools/profile_r02.sh`, after the same command had exited 0 without ncu.  Kernel sources sha (`bench.kernel_source_sha()`): `@SHA@`.\n\n```\n@SUMM@```\n\n"
docs={
 'r02_k_scaff_score.md':"# Round 2 — ncu summary of `k_scaff_score` (cfg3: 5,000 contigs / 200 M records / population 1,024)\n\n`ncu --set full --clock-control none --import-source on -k regex:k_scaff_score -s 1 -c 1 -o gpurun_out/r02_k_scaff_score python tools/prof_workload.py r3 2`\n"+HEAD+
 "Reading.  24.89 ms (27.07 ms before the record ring): every warp now takes a contiguous run of the segment's packets\nthrough its own two-slot ring of 1 KB bulk copies (3.15 M UBLKCP) instead of a one-word global load per packet with the\nnext one in a register and an L2 prefetch; 2.17e10 warp instructions = 13.6 per record-warp of 128 evaluations (17.1\nbefore), issue slots 75.8 %, **L1/shared data pipe 98.2 % busy** - the roof DESIGN.md names (4.375 wavefronts per\nrecord-warp: 7.08e9 shared wavefronts for 1.6e9 record-warps = 4.43).  DRAM: 848.0 MB read + 4.5 MB written = 2.03 x the\nalgorithmic 420.8 MB (the 8 candidate groups stream the records in two waves of four and share them through L2); this is\nwhat `profiles/r02_traffic.json` quotes.  Around the kernel: the placement tables come from `k_scaff_inv16` +\n`k_scaff_tables16` (coalesced; see the launch list), and populations that are not a multiple of 128 no longer pay for a\npadded group.\n",
 'r02_k_scaffg2_score.md':"# Round 2 — ncu summary of `k_scaffg2_score` (G-profile at cfg3 size: 5,000 contigs / 200 M records with offsets / population 1,024)\n\n`ncu --set full --clock-control none --import-source on -k regex:k_scaffg2_score -s 1 -c 1 -o gpurun_out/r02_k_scaffg2_score python tools/prof_workload.py g3 2`\n"+HEAD+
 "Reading.  119.97 ms per launch (round 1, `k_scaffg_score`: 169 ms): 8.44e10 warp instructions for 1.6e9 record-warps\n(200 M records x 8 groups of 128 candidates) = 52.7 per record-warp = **13.2 warp instructions per record per 32\ncandidates** (round 1: 23.5).  Issue slots 61 %, ALU pipe 59 %, FMA pipe 20 % (IMAD.HI counts one issue but holds the pipe\ntwice as long: measured 1.00 warp instructions/cycle/SM against 1.96 for VABSDIFF and 1.99 for IMAD, `tools/micro/alu_rate.cu`),\nL1/shared pipe 49 %, 4 warps per scheduler (128 registers/thread).  DRAM 4.12 GB read = 2.06 x the algorithmic 2.0 GB\n(records + offsets are 10 B each; two waves of four groups share them through L2).  First version of the kernel in this\nround (register rotation between three packets, post-shift on the ALU pipe): 126.5 ms, 60.9 instructions per\nrecord-warp, stalls per issue: math_pipe_throttle 1.25, not_selected 1.48, wait 0.95, dispatch 0.49.  Ceiling of the\nformulation (DESIGN.md §4): 44 ALU-pipe cycles per record-warp and scheduler -> 3.4e12 evaluations/s; achieved\n1.67e12 = 0.49.\n",
 'r02_k_scaff_stream1.md':"# Round 2 — ncu summary of `k_scaff_stream1` (one ordering against cfg3's 200 M records: the HBM-bound regime)\n\n`ncu --set full --clock-control none --import-source on -k regex:k_scaff_stream1 -s 3 -c 1 -o gpurun_out/r02_k_scaff_stream1 python tools/prof_workload.py s1 5`\n"+HEAD+
 "Reading.  75.7 us under ncu (bench.py's live events: 77.8 us = 0.787 of the measured 6.53 TB/s; 81.9 us / 0.75 with the\nper-lane cp.async ring), DRAM read 400.2 MB = the algorithmic bytes (nothing read twice).  The records now arrive by\n204,521 bulk copies of 2 KB (one per warp and ring slot): shared-memory wavefronts of the LSU pipe 16.24 M -> 14.02 M\n(20.8 -> 17.9 per warp-packet: the LDGSTS writes are gone), 6.23e7 warp instructions (6.58e7), registers 54.  What binds\nnow is the ALU pipe (80 % busy, issue slots 80 %): one LOP3 + one VIADDMNMX per record and the address extraction.\nPer-CTA time stamps (`tools/s1_trace.py`): prologue 3.5 us, stream phase 64 us per SM, ~6 us between the events and the\nfirst / after the last CTA.\n",
 'r02_k_scaff_stream1p_p2.md':"# Round 2 — ncu summary of `k_scaff_stream1p<2>` (two orderings per pass over cfg3's 200 M records)\n\n`S1_P=2 ncu --set full --clock-control none --import-source on -k regex:k_scaff_stream1p -s 3 -c 1 -o gpurun_out/r02_k_scaff_stream1p_p2 python tools/prof_workload.py s1 5`\n"+HEAD+
 "Reading.  83.2 us under ncu (bench.py: 84.0 us) for 400.1 MB -> 4.76 TB/s = 0.73 of the measured HBM peak (VERDICT r1\nasked for 0.70; 96.2 us / 0.64 with the per-lane cp.async ring): 6.91e7 warp instructions (7.86e7), L1/shared pipe 80 %,\nissue slots 79 %, ALU pipe 71 %.  The second ordering costs one more table look-up per record and the packed decay\nlook-up (23.2 shared wavefronts per warp-packet against 17.9 at P = 1).  Prologue 8.3 -> 5.5 us (both orderings' genes\nrequested ahead of the ring, pair table built at compile time, the permutation check's gathers left in flight).\n",
}
for fn,txt in docs.items():
    rep='gpurun_out/'+fn.replace('.md','.ncu-rep')
    open('profiles/'+fn,'w').write(txt.replace('@SHA@',sha).replace('@SUMM@',summ(rep)))
raw=subprocess.run(['ncu','-i','gpurun_out/r02_k_scaff_score.ncu-rep','--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(raw))); hdr,units,vals=rows[0],rows[1],rows[2]
def get(name):
    i=hdr.index(name); v=float(vals[i].replace(',','')); u=units[i]
    return v*{'Gbyte':1e9,'Mbyte':1e6,'Kbyte':1e3,'byte':1}[u]
tb=int(get('dram__bytes_read.sum')+get('dram__bytes_write.sum'))
json.dump({"_what":"dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from one ncu --set full capture; bench.py quotes it as roofline.traffic only while kernel_source_sha matches the sources the library is built from",
 "cfg3":{"kernel":"k_scaff_score","population_per_gpu":1024,"bytes":tb,"kernel_source_sha":sha,
         "source":"profiles/r02_k_scaff_score.md (ncu --set full, tools/prof_workload.py r3)",
         "note":"1.96 x the algorithmic bytes: the 8 candidate groups stream the records as two waves of 4 groups; within a wave the other three groups hit L2"}},
 open('profiles/r02_traffic.json','w'),indent=1)
print(sha,tb)
