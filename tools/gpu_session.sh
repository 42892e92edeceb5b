#!/bin/bash
# One GPU session.  usage: TAG=r3x tools/gpu_session.sh [stage ...]   (every stage writes gpurun_out/${TAG}_*; plain runs before any ncu)
#   box       machine facts                      tests     the whole -m gpu suite
#   newtests  only this round's new GPU tests    bench / ref   bench.py, our arm / the reference arm
#   ncu, ncu2 launch list + --set full capture of the inflate kernels (ncu2: COLS=1 adds the column kernels, OLD=1 resolve3)
#   multi     NG GPUs: torchrun bench + one extension query over NG GPUs        extprof   where the extension's scan threads spend their time
#   A/B stages (same process, same bytes, one switch): ab_resolve (resolve4), ab_resolve5, r5_pf, sm_reserve, ab_overlap
#   text, deflate, cfg4: the GPU tests / the at-size run of those paths
mkdir -p gpurun_out
T=${TAG:-r2}
for stage in "$@"; do
  case $stage in
    box) (nproc; free -g; lscpu | grep -E "Model name|Socket|NUMA|Thread|Core"; nvidia-smi -L) > gpurun_out/${T}_box.txt 2>&1 ;;
    smoke) timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/${T}_smoke.log ;;
    newtests) timeout 1500 python -m pytest tests/test_wire_gpu.py tests/test_ext_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_new.log 2>&1; echo "pytest(new) exit $?" | tee -a gpurun_out/${T}_pytest_new.log; tail -25 gpurun_out/${T}_pytest_new.log ;;
    tests) timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${T}_pytest_gpu.log; tail -25 gpurun_out/${T}_pytest_gpu.log ;;
    bench) timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; cat gpurun_out/${T}_bench.json; tail -5 gpurun_out/${T}_bench.err ;;
    ref) timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_reference_bench.json 2> gpurun_out/${T}_reference_bench.err; cat gpurun_out/${T}_reference_bench.json; tail -3 gpurun_out/${T}_reference_bench.err ;;
    ncu) bash tools/gpu_ncu1.sh ;;
    multi) # N GPUs (gpurun --gpus N; NG=N): bench.py under torchrun (weak + strong scaling lines), then ONE query over N GPUs through the extension
      NG=${NG:-2}
      timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $NG --steps 2 --warmup 2 --bcf-tiles 10 \
        > gpurun_out/${T}_bench_n${NG}.json 2> gpurun_out/${T}_bench_n${NG}.err; echo "bench n=$NG exit $?"; cat gpurun_out/${T}_bench_n${NG}.json; tail -3 gpurun_out/${T}_bench_n${NG}.err
      python tools/make_tiled.py /tmp/duckhts_b200_cache/tiled100m.bam 100e6 > gpurun_out/${T}_multi_ext.log 2>&1
      DEVS=$(python -c "print(','.join(str(i) for i in range($NG)))")
      for devs in 0 $DEVS; do
        DUCKHTS_DEVICES=$devs DUCKHTS_PROFILE=1 harness/_bin/fake_duckdb duckhts_b200/lib/duckhts.duckdb_extension.so read_bam /tmp/duckhts_b200_cache/tiled100m.bam standard_tags=true --cols 0,1,2,3,4,5,6,7,8,9,10,11,12,47,40,14 --threads 16 --time --repeat 3 2>&1 | sed "s/^/[devices $devs] /"
        DUCKHTS_DEVICES=$devs harness/_bin/fake_duckdb duckhts_b200/lib/duckhts.duckdb_extension.so read_bam /tmp/duckhts_b200_cache/tiled100m.bam standard_tags=true --cols 0,1,3,9,47 --threads 1 --hash 2>&1 | sed "s/^/[devices $devs, ordered hash] /"
      done >> gpurun_out/${T}_multi_ext.log 2>&1
      rm -f /tmp/duckhts_b200_cache/tiled100m.bam; cat gpurun_out/${T}_multi_ext.log
      DUCKHTS_TEST_DEVICES=$DEVS timeout 900 python -m pytest tests/test_ext_gpu.py -m gpu -q -k "several_pipelines or early_range or worker_threads" > gpurun_out/${T}_pytest_multi.log 2>&1; tail -3 gpurun_out/${T}_pytest_multi.log ;;
    extprof) # where the scan threads of the extension spend their time (DUCKHTS_PROFILE), on the bench unit, 1 / 4 / 16 threads, both layouts
      U=$(python -c "import bench,os; print(bench.gen_unit('bam', 4000000, 0x9E3779B97F4A7C15, True, os.cpu_count()))")
      for t in 1 4 16; do for lay in compact duckdb; do
        DUCKHTS_LAYOUT=$lay DUCKHTS_PROFILE=1 harness/_bin/fake_duckdb duckhts_b200/lib/duckhts.duckdb_extension.so read_bam $U standard_tags=true --cols 0,1,2,3,4,5,6,7,8,9,10,11,12,47,40,14 --threads $t --time 2>&1 | sed "s/^/[$lay t=$t] /"
      done; done > gpurun_out/${T}_extprof.log 2>&1
      python tools/make_tiled.py /tmp/duckhts_b200_cache/tiled100m.bam 100e6 >> gpurun_out/${T}_extprof.log 2>&1
      for t in 1 16; do
        DUCKHTS_PROFILE=1 harness/_bin/fake_duckdb duckhts_b200/lib/duckhts.duckdb_extension.so read_bam /tmp/duckhts_b200_cache/tiled100m.bam standard_tags=true --cols 0,1,2,3,4,5,6,7,8,9,10,11,12,47,40,14 --threads $t --time --repeat 3 2>&1 | sed "s/^/[100M t=$t] /"
      done >> gpurun_out/${T}_extprof.log 2>&1
      DH_TRACE=1 DUCKHTS_PROFILE=1 harness/_bin/fake_duckdb duckhts_b200/lib/duckhts.duckdb_extension.so read_bam /tmp/duckhts_b200_cache/tiled100m.bam standard_tags=true --cols 0,1,2,3,4,5,6,7,8,9,10,11,12,47,40,14 --threads 16 --time > gpurun_out/${T}_exttrace.log 2>&1
      rm -f /tmp/duckhts_b200_cache/tiled100m.bam
      cat gpurun_out/${T}_extprof.log; tail -5 gpurun_out/${T}_exttrace.log ;;
    ab_resolve) # inflate phase 2: the shared-memory kernel (DH_RESOLVE4=1) against the warp-per-block kernel in global memory (default)
      for v in 3 4; do
        if [ $v = 4 ]; then export DH_RESOLVE4=1; else unset DH_RESOLVE4; fi
        timeout 600 python bench.py --steps 3 --warmup 3 --no-ext --bcf-sites 0 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('resolve$v', json.dumps({'value': d['value'], 'inflate': d['inflate'], 'e2e': d['e2e']['value']}))"
      done > gpurun_out/${T}_ab_resolve.log 2>&1; unset DH_RESOLVE4; cat gpurun_out/${T}_ab_resolve.log ;;
    ab_resolve5) # inflate phase 2: the byte-parallel ring kernel (DH_RESOLVE5=1) against the warp-per-block kernel in global memory
      DH_RESOLVE5=1 timeout 240 python -m pytest tests/test_bgzf_gpu.py tests/test_zz_no_alloc.py tests/test_bam_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_resolve5.log 2>&1; echo "pytest(resolve5) exit $?" | tee -a gpurun_out/${T}_pytest_resolve5.log; tail -15 gpurun_out/${T}_pytest_resolve5.log
      for v in 0 1; do
        export DH_RESOLVE5=$v
        timeout 300 python bench.py --steps 3 --warmup 3 --no-ext --bcf-sites 0 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('resolve5=$v', json.dumps({'value': d['value'], 'ms_per_step': d['ms_per_step'], 'inflate': d['inflate'], 'kernel_ms': d['kernel_ms_per_step'], 'e2e': d['e2e']['value'], 'parity': [c['ok'] for c in d['parity']['checks']]}))"
      done > gpurun_out/${T}_ab_resolve5.log 2>&1; unset DH_RESOLVE5; cat gpurun_out/${T}_ab_resolve5.log ;;
    r5_pf) # phase 2: prefetch of old (global-memory) match sources when their descriptor is built: 0 none, 1 into L2, 2 into L1
      for v in ${PFS:-0 1 2}; do
        DH_R5_PF=$v timeout 170 python bench.py --steps 3 --warmup 3 --no-ext --bcf-sites 0 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('r5_pf=$v', json.dumps({'value': d['value'], 'ms_per_step': d['ms_per_step'], 'phase_ms': d['inflate']['phase_ms'], 'kernel_ms': d['kernel_ms_per_step'], 'parity': [c['ok'] for c in d['parity']['checks']]}))"
      done > gpurun_out/${T}_r5_pf.log 2>&1; cat gpurun_out/${T}_r5_pf.log ;;
    sm_reserve) # SMs the inflate leaves to the column kernels in device-output mode (default 32, r1k sweep); again with the new phase 2
      for v in ${RESERVES:-24 40}; do
        DH_INFLATE_SM_RESERVE=$v timeout 120 python bench.py --steps 3 --warmup 3 --no-ext --bcf-sites 0 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('reserve=$v', json.dumps({'value': d['value'], 'ms_per_step': d['ms_per_step'], 'kernel_ms': d['kernel_ms_per_step'], 'e2e': d['e2e']['value']}))"
      done > gpurun_out/${T}_sm_reserve.log 2>&1; cat gpurun_out/${T}_sm_reserve.log ;;
    ncu2) bash tools/gpu_ncu2.sh ;;
    ab_overlap) # who shares an SM with phase 1: pools of 227 KB (nobody) or 222 KB; phase 2 with / without its shared-memory tables; one or two inflate
      # streams; the other kernels asking for the same (maximal) shared-memory carve-out as phase 1 or not
      for cfg in ${AB_CFGS:-"0,0,1,0" "1,0,1,0" "1,0,1,1" "1,1,1,1" "1,1,2,1" "1,0,2,1"}; do
        IFS=, read a b c d <<< "$cfg"
        export DH_P1_SMALL=$a DH_RESOLVE_LDG=$b DH_PHASE_STREAMS=$c DH_CARVEOUT_MAX=$d
        timeout 600 python bench.py --steps 3 --warmup 3 --no-ext --bcf-sites 0 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('p1_small=$a resolve_ldg=$b phase_streams=$c carveout_max=$d', json.dumps({'value': d['value'], 'ms_per_step': d['ms_per_step'], 'inflate': d['inflate'], 'e2e': d['e2e']['value'], 'parity': d.get('parity')}))"
      done > gpurun_out/${T}_ab_overlap.log 2>&1; unset DH_P1_SMALL DH_RESOLVE_LDG DH_PHASE_STREAMS DH_CARVEOUT_MAX; cat gpurun_out/${T}_ab_overlap.log ;;
    text) timeout 1200 python -m pytest tests/test_text_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_text.log 2>&1; echo "pytest(text) exit $?" | tee -a gpurun_out/${T}_pytest_text.log; tail -40 gpurun_out/${T}_pytest_text.log ;;
    deflate) timeout 1200 python -m pytest tests/test_deflate_gpu.py -m gpu -x -q -s > gpurun_out/${T}_pytest_deflate.log 2>&1; echo "pytest(deflate) exit $?" | tee -a gpurun_out/${T}_pytest_deflate.log; tail -40 gpurun_out/${T}_pytest_deflate.log ;;
    cfg4) # BASELINE configs[3] at size: a sorted 100 M-record file, index built on the GPU and by htslib, every cell hashed against the reference
      df -h /tmp | tail -1
      timeout 2400 python tools/cfg4_100m.py --records ${CFG4_RECORDS:-100e6} --devices 0 --out gpurun_out/${T}_cfg4.json > gpurun_out/${T}_cfg4.log 2>&1; echo "cfg4 exit $?" | tee -a gpurun_out/${T}_cfg4.log
      timeout 1200 python tools/cfg4_100m.py --records ${CFG4_RECORDS:-100e6} --devices 0,0 --ordered --skip-index --out gpurun_out/${T}_cfg4_two_sources.json > gpurun_out/${T}_cfg4_two_sources.log 2>&1; echo "cfg4 (two sources, ordered) exit $?" | tee -a gpurun_out/${T}_cfg4_two_sources.log
      grep -h "cfg4\]" gpurun_out/${T}_cfg4.log gpurun_out/${T}_cfg4_two_sources.log; tail -3 gpurun_out/${T}_cfg4.log ;;
    *) echo "unknown stage $stage" ;;
  esac
done
