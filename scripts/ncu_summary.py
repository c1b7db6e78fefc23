"""Per-kernel summary (averages over the captured launches) of an `ncu --set full` raw page:
   ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python scripts/ncu_summary.py raw.csv out.json "<source note>" """
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
KEYS = {
    'gpu_time_us': 'gpu__time_duration.sum', 'dram_read_mb': 'dram__bytes_read.sum', 'dram_write_mb': 'dram__bytes_write.sum',
    'sm_throughput_pct': 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'tensor_pipe_active_pct': 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active',
    'issue_active_pct': 'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'l1tex_throughput_pct': 'l1tex__throughput.avg.pct_of_peak_sustained_active',
    'lts_throughput_pct': 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2_hit_rate_pct': 'lts__t_sector_hit_rate.pct',
    'warps_active_pct': 'sm__warps_active.avg.pct_of_peak_sustained_active', 'registers_per_thread': 'launch__registers_per_thread',
}
SCALE = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3}
S = 4 * 256 * 256 * 128      # C2: elements of the pair tensor
ALG = {'trimul_proj_tc_kernel': 10 * S, 'trimul_contract_pers_tc_kernel': 6 * S, 'trimul_contract_tc_kernel': 6 * S,
       'trimul_out_tc_kernel': 12 * S, 'pair_transition_tc_kernel': 8 * S, 'ipa_attention_tc_kernel': 2 * S,
       'pair_embed_tc_kernel': 4 * S, 'pair_embed_kernel': 4 * S, 'ipa_prep_tc_kernel': 4 * S + 2 * S + 8 * 4 * 256 * 12 * 256 * 4,
       # GB_PREC_BF16X3: hi | lo operand images, fp32 g
       'trimul_proj_x3_kernel': 10 * S, 'trimul_contract_pers_tc_kernel<1>': 12 * S, 'trimul_out_x3_kernel': 16 * S,
       'pair_transition_x3_kernel': 8 * S}


def val(r, m):
    try:
        v = float(r[idx[m]].replace(',', ''))
    except (ValueError, KeyError):
        return None
    return v * SCALE.get(units[idx[m]], 1.0)


agg = {}
for r in data:
    agg.setdefault(r[idx['Kernel Name']].split('(')[0].replace('void ', ''), []).append(r)
out = {'source': sys.argv[3] if len(sys.argv) > 3 else sys.argv[1]}
for name, rs in agg.items():
    d = {'launches_captured': len(rs)}
    for k, m in KEYS.items():
        vs = [v for v in (val(r, m) for r in rs) if v is not None]
        d[k] = round(sum(vs) / len(vs), 2) if vs else None
    d['dram_bytes_per_launch'] = round((d['dram_read_mb'] + d['dram_write_mb']) * 1e6)
    d['dram_gbs'] = round(d['dram_bytes_per_launch'] / d['gpu_time_us'] / 1e3, 1)
    for a, v in ALG.items():
        if name.split('<')[0] == a or name.replace('(bool)', '') == a:
            d['algorithmic_bytes_per_launch'] = v
    out[name] = d
json.dump(out, open(sys.argv[2], 'w'), indent=1)
for k, d in out.items():
    if isinstance(d, dict):
        print('%-36s n=%2d  %7.1f us  dram %6.1f MB %6.1f GB/s  tensor %5.1f%%  issue %5.1f%%  l1 %5.1f%%' % (
            k[:36], d['launches_captured'], d['gpu_time_us'], d['dram_bytes_per_launch'] / 1e6, d['dram_gbs'], d['tensor_pipe_active_pct'] or 0,
            d['issue_active_pct'] or 0, d['l1tex_throughput_pct'] or 0))
