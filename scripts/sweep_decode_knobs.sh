run() { name=$1; shift; env "$@" python bench.py --steps 64 --warmup 8 --no-ttft 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$name', d['value'], d['ms_per_step'])"; }
run base A=1
run pf_32_16_16_16 B200_L2_PREFETCH_MB=32,16,16,16
run pf_16_16_16_16 B200_L2_PREFETCH_MB=16,16,16,16
run pf_24_24_24_24 B200_L2_PREFETCH_MB=24,24,24,24
run pf_24_8_16_8 B200_L2_PREFETCH_MB=24,8,16,8
run splits4 B200_ATTN_SPLITS=4
run splits8 B200_ATTN_SPLITS=8
run bulk0 B200_ATTN_BULK=0
