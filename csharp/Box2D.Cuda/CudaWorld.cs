// Hand-written facade over Native.cs: how the reference's static Box2D.API surface routes to libb2cuda.so. SOURCE ONLY — no .NET toolchain
// exists in the build image, so this file is not compiled there; the identical call sequence is exercised through ctypes
// (box2d_b200/api.py) and checked by tests/. Reference lines: api/World.cs:12-113,283-319; api/Body.cs:16-160; api/Shape.cs:8-56; api/Joint.cs.
using System;
using System.Runtime.InteropServices;
namespace Box2D.Cuda;

public sealed unsafe class CudaWorld : IDisposable
{
    nint handle;
    static void Check(int rc)
    {
        if (rc == 0) return;
        string msg = Marshal.PtrToStringUTF8((nint)Native.b2cuda_last_error()) ?? "";
        throw rc == -2 ? new NotSupportedException(msg) : new InvalidOperationException($"libb2cuda rc={rc}: {msg}");   // B2C_E_UNSUPPORTED = -2
    }
    public CudaWorld(in B2CWorldDef def, int device = 0)                    // WorldAPI.CreateWorld api/World.cs:12-16
    {
        fixed (B2CWorldDef* d = &def) { nint h; Check(Native.b2cuda_world_create(d, device, &h)); handle = h; }
    }
    public void Dispose() { if (handle != 0) { Native.b2cuda_world_destroy(handle); handle = 0; } }   // WorldAPI.DestroyWorld :18
    public void Step(float timeStep, int subStepCount) => Check(Native.b2cuda_world_step(handle, timeStep, subStepCount));   // WorldAPI.Step :31-113 (asynchronous)
    public int CreateBody(in B2CBodyDef def)                                // BodyAPI.CreateBody api/Body.cs:16-105; the C# side wraps the int as BodyID{index1 = id + 1}
    {
        fixed (B2CBodyDef* d = &def) { int id; Check(Native.b2cuda_create_body(handle, d, &id)); return id; }
    }
    public int CreateShape(int bodyId, in B2CShapeDef def, in B2CGeometry geometry)   // ShapeAPI.CreateShape api/Shape.cs:8-24 (geometry flattened from Geometry.MakeBox etc.)
    {
        fixed (B2CShapeDef* d = &def) fixed (B2CGeometry* g = &geometry) { int id; Check(Native.b2cuda_create_shape(handle, bodyId, d, g, &id)); return id; }
    }
    public int CreateJoint(in B2CJointDef def)                              // JointAPI.Create*Joint api/Joint.cs:182,324,408,412,555,672,720
    {
        fixed (B2CJointDef* d = &def) { int id; Check(Native.b2cuda_create_joint(handle, d, &id)); return id; }
    }
    public void DestroyBody(int bodyId) => Check(Native.b2cuda_destroy_body(handle, bodyId));                     // api/Body.cs:106-160
    public void DestroyShape(int shapeId, bool updateBodyMass) => Check(Native.b2cuda_destroy_shape(handle, shapeId, updateBodyMass ? 1 : 0));   // api/Shape.cs:48-56
    public void DestroyJoint(int jointId, bool wakeAttached) => Check(Native.b2cuda_destroy_joint(handle, jointId, wakeAttached ? 1 : 0));     // api/Joint.cs:9-17

    /// WorldAPI.GetContactEvents api/World.cs:302-313: spans over library-owned pinned memory, valid until the next Step ("transient" contract :301)
    public void GetContactEvents(out ReadOnlySpan<B2CContactTouchEvent> begin, out ReadOnlySpan<B2CContactTouchEvent> end, out ReadOnlySpan<B2CContactHitEvent> hit)
    {
        B2CContactTouchEvent* b, e; B2CContactHitEvent* h; int nb, ne, nh;
        Check(Native.b2cuda_world_get_contact_events(handle, &b, &nb, &e, &ne, &h, &nh));
        begin = new(b, nb); end = new(e, ne); hit = new(h, nh);
    }
    /// WorldAPI.GetSensorEvents api/World.cs:289-299
    public void GetSensorEvents(out ReadOnlySpan<B2CSensorEvent> begin, out ReadOnlySpan<B2CSensorEvent> end)
    {
        B2CSensorEvent* b, e; int nb, ne;
        Check(Native.b2cuda_world_get_sensor_events(handle, &b, &nb, &e, &ne));
        begin = new(b, nb); end = new(e, ne);
    }
    /// WorldAPI.GetBodyEvents api/World.cs:283-287: one move event per awake body (transform, body id, fellAsleep)
    public ReadOnlySpan<B2CBodyMoveEvent> GetBodyMoveEvents()
    {
        B2CBodyMoveEvent* p; int n;
        Check(Native.b2cuda_world_get_body_move_events(handle, &p, &n));
        return new(p, n);
    }
    /// bulk form of BodyAPI.GetTransform / GetLinearVelocity / GetAngularVelocity (api/Body.cs:282-357)
    public int ReadBodies(Span<B2CBodyRead> dst)
    {
        fixed (B2CBodyRead* p = dst) { int n; Check(Native.b2cuda_world_read_bodies(handle, p, dst.Length, &n)); return n; }
    }
}
