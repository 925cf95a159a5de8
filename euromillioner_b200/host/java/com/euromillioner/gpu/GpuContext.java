package com.euromillioner.gpu;

import java.util.HashMap;
import java.util.Map;

/**
 * One engine context per GPU and process, shared by every estimator and model of that device: a context owns the
 * fit-sized device buffers, so creating one per fit() (as round 1's estimator did) re-allocated gigabytes per call.
 * Reference counted; the native context is destroyed when the last user closes.  SOURCE ONLY.
 */
public final class GpuContext implements AutoCloseable {
    private static final Map<Integer, GpuContext> SHARED = new HashMap<>();
    final long handle;
    private final Integer sharedKey;   // null: a distributed context (one per rank, never shared)
    private int refs = 1;

    private GpuContext(long handle, Integer sharedKey) { this.handle = handle; this.sharedKey = sharedKey; }

    /** The process-wide context of `device`, created on first use. */
    public static GpuContext forDevice(int device) {
        synchronized (SHARED) {
            GpuContext c = SHARED.get(device);
            if (c == null) { c = new GpuContext(Native.ctxCreate(device), device); SHARED.put(device, c); }
            else c.refs++;
            return c;
        }
    }
    /** Multi-GPU: rank `rank` of `world`, with the id rank 0 obtained from ncclUniqueId() and sent to the others. */
    public static GpuContext distributed(int device, int rank, int world, byte[] ncclUniqueId) {
        return new GpuContext(Native.ctxCreateDist(device, rank, world, ncclUniqueId), null);
    }
    public static byte[] ncclUniqueId() { return Native.ncclUniqueId(); }

    GpuContext retain() { synchronized (SHARED) { refs++; } return this; }
    /** Between jobs: device buffers and cached blocks go back to the driver (the next fit re-allocates). */
    public long releaseBuffers() { return Native.ctxReleaseBuffers(handle); }

    @Override public void close() {
        synchronized (SHARED) {
            if (--refs > 0) return;
            if (sharedKey != null) SHARED.remove(sharedKey);
        }
        Native.ctxDestroy(handle);
    }
}
