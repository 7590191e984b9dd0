timeout 200 python tools/knob_sweep.py 8,20,-1,-1,-1,-1,5 8,20,-1,-1,-1,-1,1 8,20,-1,-1,-1,-1,5 8,20,-1,-1,-1,-1,1 2>&1 | tail -5
bash tools/probe/quick.sh 2>&1 | cut -c1-330
