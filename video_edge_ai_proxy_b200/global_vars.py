"""Shared producer/consumer state — same names as reference python/global_vars.py:15-17.

The reference imports `query_timestamp` BY VALUE into two modules (`from global_vars import
query_timestamp`, read_image.py:20 and rtsp_to_rtmp.py:28), so the consumer never sees the
producer's updates.  Users of this module must read and write the attribute on the
module (`global_vars.query_timestamp = ts`), which is what `read_image.ReadImage` here does.
"""
query_timestamp = None
RedisLastAccessPrefix = "last_access_time_"
RedisIsKeyFrameOnlyPrefix = "is_key_frame_only_"
