"""CPU tests of the profiling helpers under tools/ (no GPU, no native library)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CSV = '''==PROF== Connected to process 1
"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size","Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"
"0","1","python","h","void pmd::<unnamed>::pad_stem_weight_kernel(const __nv_bfloat16 *, __nv_bfloat16 *, int, int)","1","14","(256, 1, 1)","(16, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","2900"
"1","1","python","h","void pmd::<unnamed>::augment_im2col_kernel(const unsigned char *, const int *, int, int, int, int, __nv_bfloat16 *)","1","14","(128, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","31000"
"2","1","python","h","void pmd::<unnamed>::conv_gemm_kernel<64, 2, 0, 0>(pmd::ConvGemmParams)","1","14","(192, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","48,000"
"3","1","python","h","void pmd::<unnamed>::conv_gemm_kernel<64, 2, 0, 0>(pmd::ConvGemmParams)","1","14","(192, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","52000"
"4","1","python","h","void pmd::<unnamed>::pad_stem_weight_kernel(const __nv_bfloat16 *, __nv_bfloat16 *, int, int)","1","14","(256, 1, 1)","(16, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","3000"
"5","1","python","h","void pmd::<unnamed>::augment_im2col_kernel(const unsigned char *, const int *, int, int, int, int, __nv_bfloat16 *)","1","14","(128, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","30000"
'''


def test_launch_table_picks_the_last_complete_step(tmp_path):
    src = tmp_path / 'l.csv'
    src.write_text(CSV)
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'launch_table.py'), str(src), 'unit test'],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    assert out.startswith('# unit test')
    # the step is launches 1..4 (the first kernel of a step .. just before the next step's first kernel)
    assert '| conv_gemm_kernel<64, 2, 0, 0> | 2 | 100.0 |' in out
    assert '| **total (serialised)** | 4 | 134.0 | |' in out
    assert '| 0 | augment_im2col_kernel | (4096, 1, 1) | 31.0 |' in out


def test_ncu_table_aggregates_time_weighted_tensor_activity(tmp_path):
    cols = ['ID', 'Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'launch__registers_per_thread',
            'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
            'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
            'lts__throughput.avg.pct_of_peak_sustained_elapsed']
    units = ['', '', '', 'us', 'register/thread', '%', '%', '%', 'Mbyte', 'Kbyte', '%']
    rows = [['0', 'void unnamed>::conv_gemm_kernel<128, 3, 0, 0, 0>(pmd::ConvGemmParams)', '(1024, 1, 1)', '50.0', '56', '17', '40.0',
             '10', '67.0', '500', '30'],
            ['1', 'void unnamed>::conv_gemm_kernel<128, 3, 0, 0, 0>(pmd::ConvGemmParams)', '(1024, 1, 1)', '150.0', '56', '17', '20.0',
             '10', '67.0', '500', '30'],
            ['2', 'void unnamed>::bn_finalize_apply_kernel<0>(int)', '(592, 1, 1)', '100.0', '40', '60', '0.0', '60', '134.0',
             '67000', '40']]
    src = tmp_path / 'raw.csv'
    src.write_text('\n'.join(','.join('"%s"' % c for c in r) for r in [cols, units] + rows) + '\n')
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'ncu_table.py'), str(src), 'unit test'],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    assert '| 0 | conv_gemm_kernel<128, 3, 0, 0, 0> | (1024, 1, 1) | 50.0 | 56 | 17.0 | 40.0 | 10.0 | 67.0 | 0.5 | 30.0 |' in out
    assert '| conv_gemm_kernel<128, 3, 0, 0, 0> | 2 | 200.0 | 25.0 | 10.0 |' in out          # (50*40 + 150*20) / 200
    assert '| **all 3 launches** | | 300.0 | 16.7 | |' in out
    assert '| **convolution kernels only** | | 200.0 | 25.0 | |' in out
