export SHAPES="1024,32,1,1,272,480;1024,48,1,1,272,480;1024,64,1,1,272,480;1024,96,1,1,272,480;1024,128,1,1,272,480;1024,192,1,1,272,480;1024,256,1,1,272,480"
echo "DEBUG=5 (no TMA, no stores): MMA + handshakes only"
DCVC_TC_DEBUG=5 python scripts/bench_conv.py 1 | cut -c1-40,60-100
echo "DEBUG=7 (no MMA either)"
DCVC_TC_DEBUG=7 python scripts/bench_conv.py 1 | cut -c1-40,60-100
echo "DEBUG=5 CTAS=1"
DCVC_TC_CTAS=1 DCVC_TC_DEBUG=5 python scripts/bench_conv.py 1 | cut -c1-40,60-100
