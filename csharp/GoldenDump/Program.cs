// Golden-vector dump for the REFERENCE itself (MWstudios/Box2D, C#/.NET 9). SOURCE ONLY: the build image has no dotnet / mono / csc (probed by
// bench.py at run time), so this has never been run here; it is what a maintainer with a .NET 9 SDK runs once to PIN the oracle:
//   dotnet new console -o GoldenDump && copy this file over Program.cs && dotnet add reference ../../<reference>/Box2D.csproj && dotnet run -c Release -- out/
// It builds the scenes of box2d_b200/scenes.py through Box2D.API (same calls, same order, same fp32 literals), steps them with dt = 1/60 and
// 4 substeps on one thread, and writes per step the quantities tests/golden/make_golden.py records from the oracle: contact keys in
// contact-id order with colour, manifold point count and feature ids; begin / end touch events; body transforms and velocities as raw fp32 bits.
// tests/golden/*.npz hold the same fields from oracle/; comparing the two files pins (or refutes) the restatement.
using System;
using System.IO;
using Box2D;
using Box2D.API;

static class GoldenDump
{
    static uint Bits(float f) => BitConverter.SingleToUInt32Bits(f);

    static WorldID MakeWorld()
    {
        WorldDef def = Types.DefaultWorldDef();
        def.enableSleep = false;                   // the CUDA library refuses sleep; the goldens are recorded without it
        return WorldAPI.CreateWorld(ref def);
    }

    // scenes.add_pyramid(world, rows): static ground box with its top at y = 0, rows of unit boxes, row i has rows - i boxes
    static void Pyramid(WorldID world, int rows)
    {
        BodyDef bd = Types.DefaultBodyDef();
        BodyID ground = BodyAPI.CreateBody(world, ref bd);
        ShapeDef sd = Types.DefaultShapeDef();
        sd.enableContactEvents = true;
        Polygon groundBox = Geometry.MakeOffsetBox(rows, 1.0f, new Vector2(0.0f, -1.0f), Rotation.Identity);
        ShapeAPI.CreatePolygonShape(ground, ref sd, ref groundBox);
        Polygon box = Geometry.MakeBox(0.5f, 0.5f);
        for (int i = 0; i < rows; i++)
        {
            int count = rows - i;
            for (int j = 0; j < count; j++)
            {
                BodyDef d = Types.DefaultBodyDef();
                d.type = BodyType.Dynamic;
                d.position = new Vector2((j - (count - 1) * 0.5f) * 1.0f, 0.5f + i * 1.0f);
                BodyID b = BodyAPI.CreateBody(world, ref d);
                ShapeAPI.CreatePolygonShape(b, ref sd, ref box);
            }
        }
    }

    static void Dump(string path, WorldID worldId, int steps)
    {
        using var w = new StreamWriter(path);
        World world = worldId.index1;
        for (int step = 0; step < steps; step++)
        {
            WorldAPI.Step(worldId, 1.0f / 60.0f, 4);
            ContactEvents ev = WorldAPI.GetContactEvents(worldId);
            w.WriteLine($"step {step} begin {ev.beginEvents.Count} end {ev.endEvents.Count} hit {ev.hitEvents.Count}");
            foreach (var e in ev.beginEvents) w.WriteLine($"B {e.shapeIdA.index1 - 1} {e.shapeIdB.index1 - 1} {e.contactId.index1 - 1}");
            foreach (var e in ev.endEvents) w.WriteLine($"E {e.shapeIdA.index1 - 1} {e.shapeIdB.index1 - 1} {e.contactId.index1 - 1}");
            for (int id = 0; id < world.contacts.Count; id++)
            {
                Contact c = world.contacts[id];
                if (c.contactId != id) continue;
                ContactSim sim = world.GetContactSim(c);
                Manifold m = sim.manifold;
                w.WriteLine($"C {id} {c.shapeIdA} {c.shapeIdB} color {c.colorIndex} points {m.pointCount} ids {m.point0.id} {m.point1.id} " +
                            $"n {Bits(m.normal.x):x8} {Bits(m.normal.y):x8} sep {Bits(m.point0.separation):x8} {Bits(m.point1.separation):x8}");
            }
            for (int id = 0; id < world.bodies.Count; id++)
            {
                Body b = world.bodies[id];
                if (b.id != id) continue;
                BodySim s = world.GetBodySim(b);
                Vector2 v = BodyAPI.GetLinearVelocity(world.MakeBodyID(id));
                float av = BodyAPI.GetAngularVelocity(world.MakeBodyID(id));
                w.WriteLine($"X {id} {Bits(s.transform.p.x):x8} {Bits(s.transform.p.y):x8} {Bits(s.transform.q.c):x8} {Bits(s.transform.q.s):x8} {Bits(v.x):x8} {Bits(v.y):x8} {Bits(av):x8}");
            }
        }
    }

    static void Main(string[] args)
    {
        string outDir = args.Length > 0 ? args[0] : "golden_ref";
        Directory.CreateDirectory(outDir);
        WorldID w = MakeWorld();
        Pyramid(w, 10);
        Dump(Path.Combine(outDir, "pyramid10.txt"), w, 30);
        WorldAPI.DestroyWorld(w);
        // the other cases of tests/golden/make_golden.py (mixed_drop, joint_grid8, tumbler300, joint_zoo_exact, wheel_rig, bounce_mix) follow the
        // same pattern: transcribe the scene function of box2d_b200/scenes.py call by call, then Dump.
    }
}
