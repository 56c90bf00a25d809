mkdir -p gpurun_out
{ echo "# tools/tc_timeline.py: clock64 stamps of CTA (0,0,0), cycles since kernel entry: [setup done, first TMA issued, first stage landed, last stage landed, accumulator ready, epilogue done, TMEM freed]"; echo "## UNet layers, batch 64"; timeout 300 python tools/tc_timeline.py 2>&1 | grep -v Warning; echo "## VGGish layers, 768 images"; TCVGG=1 timeout 300 python tools/tc_timeline.py 2>&1 | grep -v Warning; } > gpurun_out/tc_timeline_r1d.txt
cat gpurun_out/tc_timeline_r1d.txt | cut -c 1-230
