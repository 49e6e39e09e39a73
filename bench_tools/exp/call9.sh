for cfg in "6 8" "4 8" "3 8" "4 4" "3 4" "6 4"; do set -- $cfg
  echo "== DPS_STAGES1=$1 DPS_WH_STAGES=$2"
  DPS_STAGES1=$1 DPS_WH_STAGES=$2 python bench.py --steps 200 --warmup 10 --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"
done
