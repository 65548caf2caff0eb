set -e
python tools/knob_rate.py
PSIFT_MARCH_MIN=2000000 python tools/knob_rate.py
PSIFT_MARCH_MIN=500000 python tools/knob_rate.py
PSIFT_MARCH_WPS=4 python tools/knob_rate.py
PSIFT_MARCH_WPS=6 python tools/knob_rate.py
PSIFT_MARCH_WPS=12 python tools/knob_rate.py
PSIFT_MARCH_MIN=2000000 PSIFT_MARCH_WPS=4 python tools/knob_rate.py
PSIFT_BLUR_SMALL_MAX=1000000 python tools/knob_rate.py
PSIFT_SLOTS=8 python tools/knob_rate.py
PSIFT_SLOTS=4 python tools/knob_rate.py
