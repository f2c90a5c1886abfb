python tools/bench_brief.py --steps 4
python tools/bench_brief.py --steps 4 SSI_B200_STORE_HINT=1
python tools/bench_brief.py --steps 4 SSI_B200_GROUP_M_GRAD=64
python tools/bench_brief.py --steps 4 SSI_B200_GROUP_M_GRAD=64 SSI_B200_STORE_HINT=1
python tools/bench_brief.py --steps 4 SSI_B200_GROUP_M_GRAD=148 SSI_B200_STORE_HINT=1
python tools/bench_brief.py --steps 4 SSI_B200_GROUP_M_DX=37 SSI_B200_GROUP_M_DW=8
python tools/bench_brief.py --steps 4 SSI_B200_G_TILES=37
python tools/bench_brief.py --steps 4 --config c3
python tools/bench_brief.py --steps 4 --config c3 SSI_B200_STORE_HINT=1 SSI_B200_GROUP_M_GRAD=64
