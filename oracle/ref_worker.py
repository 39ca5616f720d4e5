"""One worker process of the reference CPU arm: the UNMODIFIED reference package, one core.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/__init__.py).  Started by ``oracle.RefPool`` as
``python oracle/ref_worker.py``; speaks JSON lines on stdin/stdout:

    {"cmd": "load", "root": <dir holding meterelf/>, "params": <params.yml>, "frames": <.npy of
     (n, h, w, 3) u8 meter_rect crops>, "lo": a, "hi": b}         -> {"ready": b - a}
    {"cmd": "run"}    reads frames [lo, hi) through ImageFile(name, params, bgr_image=crop) and
                      get_meter_value (the reference's _reading.py:19-115)  -> {"dt": s, "results": [...]}
    {"cmd": "quit"}

The reference runs under the three environment shims of SURVEY.md 8c (tools/ref_harness.py):
yaml SafeLoader default, 3-value cv2.findContours, IPP off; cv2 is held to one thread so that one
worker = one core.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), 'tools'))


def main():
    real_out = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)                      # nothing but our JSON lines may reach the pipe
    state = {}

    def reply(obj):
        real_out.write(json.dumps(obj) + '\n')
        real_out.flush()

    for line in sys.stdin:
        msg = json.loads(line)
        if msg['cmd'] == 'quit':
            break
        if msg['cmd'] == 'load':
            os.environ['MELF_REFERENCE_ROOT'] = msg['root']
            import numpy as np
            import ref_harness
            ref_harness.REFERENCE_ROOT = msg['root']
            ref_harness.load()
            from meterelf import _image, _params, _reading, exceptions
            params = _params.load(msg['params'])
            if msg.get('threshold') is not None:
                params.dials_match_threshold = msg['threshold']
            frames = np.load(msg['frames'], mmap_mode='r')
            state.update(np=np, params=params, frames=frames, lo=msg['lo'], hi=msg['hi'],
                         image=_image, reading=_reading, exc=exceptions,
                         names=list(params.dial_centers.keys()))
            if msg['hi'] > msg['lo']:
                _warm(state)
            reply({'ready': msg['hi'] - msg['lo']})
            continue
        if msg['cmd'] == 'run':
            np = state['np']
            (image, reading, exc, params, names) = (state['image'], state['reading'], state['exc'],
                                                    state['params'], state['names'])
            out = []
            t0 = time.perf_counter()
            for i in range(state['lo'], state['hi']):
                crop = np.ascontiguousarray(state['frames'][i])
                imgf = image.ImageFile('frame%d' % i, params, crop)
                rec = {'i': i}
                try:
                    values = reading.get_meter_value(imgf)
                    rec['status'] = 0
                    rec['dial_pos'] = [float(values[k]).hex() for k in names]
                    rec['value'] = float(values['value']).hex() if 'value' in values else None
                except exc.DialsNotFoundError as e:
                    rec.update(status=1, max_val=float(e.extra_info['match val']))
                except exc.NeedleContoursNotFoundError as e:
                    rec.update(status=2, failed_dial=names.index(e.extra_info['dial']))
                except exc.DialAngleDeterminingError as e:
                    bad = e.extra_info['unreadable dials'].split(', ')
                    rec.update(status=3, unreadable_mask=sum(1 << names.index(b) for b in bad))
                out.append(rec)
            reply({'dt': time.perf_counter() - t0, 'results': out if msg.get('results', True) else None})


def _warm(state):
    """One untimed read (template load, dial masks, OpenCV's lazy initialisation)."""
    try:
        crop = state['np'].ascontiguousarray(state['frames'][state['lo']])
        state['reading'].get_meter_value(state['image'].ImageFile('warm', state['params'], crop))
    except Exception:
        pass


if __name__ == '__main__':
    main()
