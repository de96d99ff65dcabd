"""TEST INFRASTRUCTURE ONLY — literal restatement of the decode bookkeeping in reference
python/read_image.py:57-128 (variable names kept), with the one repair SURVEY §3.2 describes:
`query_timestamp` is the value the producer thread set (rtsp_to_rtmp.py:148), passed in with each
packet instead of being read from a module-local copy that is always None.

`run(events)` consumes (is_keyframe, is_corrupt, decode_only_keyframes, query_timestamp,
frames_in_packet) tuples — one per packet popped from the queue — and returns the list of
published frames as dicts with the VideoFrame bookkeeping fields.  No reference test pins this
logic ("parity unpinned" by the reference); tests compare the product gate against this walk.
"""


def run(events):
    packet_group = []
    packet_count = 0
    keyframes_count = 0
    last_query_timestamp = 0
    published = []
    for n, (is_keyframe, is_corrupt, decode_only_keyframes, query_timestamp, nframes) in enumerate(events):
        if is_keyframe:                                   # :63-66
            packet_group = []
            packet_count = 0
            keyframes_count = keyframes_count + 1
        packet_group.append((n, nframes))                 # :68
        should_decode = False                             # :70-74
        if query_timestamp is not None and query_timestamp > last_query_timestamp:
            should_decode = True
        if decode_only_keyframes:                         # :77-78
            should_decode = False
        if len(packet_group) == 1 or should_decode:       # :80
            for index, p in enumerate(packet_group):      # :81
                if index < packet_count:                  # :84-85
                    continue
                for f in range(p[1]):                     # :87  for frame in p.decode()
                    published.append(dict(source_packet=p[0], group_index=index, frame=f,
                                          packet=packet_count, keyframe=keyframes_count,
                                          is_keyframe=is_keyframe, is_corrupt=is_corrupt))
                    if decode_only_keyframes:             # :123-124
                        break
                    if query_timestamp is not None:       # never None in the reference at this point:
                        last_query_timestamp = query_timestamp  # :126 (packets are popped only after
                        #                                         rtsp_to_rtmp.py:146-153 set it)
                packet_count = packet_count + 1           # :128
    return published
