for cb in 16000000 32000000 48000000 96000000 192000000; do for cr in 1000000 4000000; do
NTC_MAX_CHUNK_BASES=$cb NTC_MAX_CHUNK_RECORDS=$cr python bench.py --steps 3 --cpu-budget 0 2>/dev/null | python -c "
import json,sys; b=json.loads(sys.stdin.read()); e=b['e2e']; print('chunk_bases=$cb chunk_records=$cr e2e=%.0f enc=%.2f ms dec=%.2f ms'%(e['value'],e['encode_ms_per_step'],e['decode_ms_per_step']))"
done; done
