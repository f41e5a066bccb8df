set -x
for rep in 1 2; do
(cd build/old_tree && timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-profile 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('OLD infer',d['value'],'e2e',d['e2e']['value'],'packed',d['e2e_packed']['value'],'train',d['train']['value'],'train e2e',d['train']['e2e']['value'])")
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-profile 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('NEW infer',d['value'],'e2e',d['e2e']['value'],'packed',d['e2e_packed']['value'],'train',d['train']['value'],'train e2e',d['train']['e2e']['value'])"
done
