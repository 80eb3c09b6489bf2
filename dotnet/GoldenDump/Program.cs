// GoldenDump -- runs the UNMODIFIED reference (its compiled Minimal_RAG.dll) on the inputs of tests/golden/*.json and
// writes what it returns to tests/golden/from_dotnet/, in the schema tests/golden/make_golden.py writes.
//
// SOURCE ONLY: this file has never been compiled or run where the repository was written (no dotnet toolchain there).
// It is test infrastructure; nothing in the product depends on it.
//
// Why it exists: the CPU oracle (oracle/ragindex_oracle.c) is a restatement of Helpers/RagIndex.cs:59-67,124-135, and
// nothing the reference ships pins it.  This tool closes that gap in one command on any machine with a .NET 9 SDK:
//
//     dotnet build <reference>/Minimal_Agent_RAG/Minimal_RAG                       # the reference, as it is
//     dotnet run --project dotnet/GoldenDump -- <reference>/Minimal_Agent_RAG/Minimal_RAG/bin/Debug/net9.0/Minimal_RAG.dll tests/golden
//     python -m pytest tests/test_oracle.py -k from_dotnet                          # oracle == reference, bit for bit
//
// How it stays honest: no line of the reference is copied here.  The tool loads the reference's own assembly and calls
//   * RagIndex.QueryAsync(query, topK)   (RagIndex.cs:59-67)  -- through the reference's own Embedder: QueryAsync asks
//     http://localhost:11434/api/embeddings for the query vector (Embedder.cs:9,34), so the tool answers on that port
//     with the test's query vector, every float printed as a round-trippable double ((float)GetDouble() gives it back
//     exactly, Embedder.cs:44).  Stop a real Ollama on that port first; on Windows HttpListener needs the URL reserved
//     (netsh http add urlacl url=http://localhost:11434/ user=<you>) or an elevated prompt.
//   * RagIndex.Cosine(float[], float[])  (RagIndex.cs:124-135) -- directly, by reflection, for every (query, row).
// Rows go straight into the private `_chunks` list (RagIndex.cs:8) as RagChunk records (Program.cs:326), which is
// what BuildFromFolderAsync does at :48; ids are the row numbers.
using System.Collections;
using System.Globalization;
using System.Net;
using System.Reflection;
using System.Text;
using System.Text.Json;
using System.Text.Json.Nodes;

static class GoldenDump
{
    static readonly Dictionary<string, float[]> Vectors = new();   // "embedding" the fake endpoint returns for a text

    static string R(float f) =>                                      // JSON has no NaN/Infinity: the KAT queries have none
        ((double)f).ToString("R", CultureInfo.InvariantCulture);

    static void ServeEmbeddings(HttpListener l)
    {
        while (l.IsListening)
        {
            HttpListenerContext ctx;
            try { ctx = l.GetContext(); } catch { return; }
            using var rd = new StreamReader(ctx.Request.InputStream, Encoding.UTF8);
            var body = JsonNode.Parse(rd.ReadToEnd())!;
            var key = body["input"] is JsonValue v ? v.GetValue<string>() : (body["prompt"]?.GetValue<string>() ?? "");
            var vec = Vectors[key];
            var json = "{\"embedding\":[" + string.Join(",", vec.Select(R)) + "]}";
            var bytes = Encoding.UTF8.GetBytes(json);
            ctx.Response.ContentType = "application/json";
            ctx.Response.OutputStream.Write(bytes, 0, bytes.Length);
            ctx.Response.Close();
        }
    }

    static float ToF32(JsonNode n) =>
        n is JsonValue jv && jv.TryGetValue<string>(out var s)
            ? (s == "nan" ? float.NaN : float.Parse(s, CultureInfo.InvariantCulture))
            : (float)n.GetValue<double>();

    // ---- the synthetic generator of oracle/ragindex_oracle.c (rag_synth_rows), restated; checked below against the
    // pinned values in synth_v1.json before it is used
    static ulong SplitMix(ulong z)
    {
        unchecked
        {
            z += 0x9E3779B97F4A7C15UL;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9UL;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBUL;
            return z ^ (z >> 31);
        }
    }
    static int Sample(ulong rowKey, long col)
    {
        unchecked
        {
            ulong u = SplitMix(rowKey ^ ((ulong)col * 0xD1B54A32D192ED03UL + 0x8CB92BA72F3D8DD7UL));
            return (int)(u & 0xFFFF) + (int)((u >> 16) & 0xFFFF) + (int)((u >> 32) & 0xFFFF) + (int)((u >> 48) & 0xFFFF) - 131070;
        }
    }
    static float Bf16Round(float x)
    {
        uint u = BitConverter.SingleToUInt32Bits(x);
        if ((u & 0x7F800000u) != 0x7F800000u) unchecked { u += 0x7FFFu + ((u >> 16) & 1u); }
        return BitConverter.UInt32BitsToSingle(u & 0xFFFF0000u);
    }
    static float[][] SynthRows(long row0, int n, int dim, ulong seed, bool bf16 = false)
    {
        var rows = new float[n][];
        for (int r = 0; r < n; r++)
        {
            ulong key;
            unchecked { key = SplitMix(seed + 0x632BE59BD9B4E019UL * (ulong)(row0 + r + 1)); }
            long n2 = 0;
            var s = new int[dim];
            for (int c = 0; c < dim; c++) { s[c] = Sample(key, c); n2 += (long)s[c] * s[c]; }
            double norm = Math.Sqrt((double)n2);
            rows[r] = new float[dim];
            for (int c = 0; c < dim; c++)
            {
                float f = (float)(n2 != 0 ? s[c] / norm : 0.0);
                rows[r][c] = bf16 ? Bf16Round(f) : f;
            }
        }
        return rows;
    }

    // ---- the reference, by reflection
    static Type ragIndexT = null!, ragChunkT = null!;
    static MethodInfo cosine = null!, queryAsync = null!;
    static FieldInfo chunksF = null!;

    static (List<int> idx, List<double> score) Query(float[][] rows, float[] q, int topK, string key)
    {
        var index = Activator.CreateInstance(ragIndexT, BindingFlags.Public | BindingFlags.NonPublic | BindingFlags.Instance, null,
                                             new object[] { "golden-dump" }, null)!;
        var chunks = (IList)chunksF.GetValue(index)!;
        for (int r = 0; r < rows.Length; r++)     // RagChunk(Id, Source, Text, Embedding): Id = the row number
            chunks.Add(Activator.CreateInstance(ragChunkT, r.ToString(CultureInfo.InvariantCulture), "golden", "", rows[r])!);
        Vectors[key] = q;
        var task = (Task)queryAsync.Invoke(index, new object[] { key, topK })!;
        task.GetAwaiter().GetResult();
        var hits = (IEnumerable)task.GetType().GetProperty("Result")!.GetValue(task)!;
        var idx = new List<int>();
        var sc = new List<double>();
        foreach (var h in hits)
        {
            idx.Add(int.Parse((string)h.GetType().GetProperty("Id")!.GetValue(h)!, CultureInfo.InvariantCulture));
            sc.Add((double)h.GetType().GetProperty("Score")!.GetValue(h)!);
        }
        return (idx, sc);
    }

    static string Bits(double d) => BitConverter.DoubleToInt64Bits(d).ToString("x16");

    static int Main(string[] args)
    {
        if (args.Length < 2) { Console.Error.WriteLine("usage: GoldenDump <path to the reference's Minimal_RAG.dll> <tests/golden directory>"); return 2; }
        var asm = Assembly.LoadFrom(Path.GetFullPath(args[0]));
        var gold = args[1];
        const BindingFlags any = BindingFlags.Public | BindingFlags.NonPublic | BindingFlags.Static | BindingFlags.Instance;
        ragIndexT = asm.GetType("Minimal_RAG.Helpers.RagIndex", true)!;
        ragChunkT = asm.GetType("RagChunk", true)!;            // a record of the top-level program: global namespace (Program.cs:326)
        cosine = ragIndexT.GetMethod("Cosine", any)!;
        queryAsync = ragIndexT.GetMethod("QueryAsync", any)!;
        chunksF = ragIndexT.GetField("_chunks", any)!;

        using var listener = new HttpListener();
        listener.Prefixes.Add("http://localhost:11434/");     // where the reference's Embedder looks (Embedder.cs:9)
        listener.Start();
        var server = new Thread(() => ServeEmbeddings(listener)) { IsBackground = true };
        server.Start();
        Directory.CreateDirectory(Path.Combine(gold, "from_dotnet"));

        // ---- known-answer cases: same inputs as kat_v1.json, outputs from the reference
        // Python's json module writes a bare NaN (not JSON): quote it, ToF32 reads "nan"
        var katText = System.Text.RegularExpressions.Regex.Replace(File.ReadAllText(Path.Combine(gold, "kat_v1.json")), @"(?<![\w""])NaN(?![\w""])", "\"nan\"");
        var kat = JsonNode.Parse(katText)!.AsArray();
        var katOut = new JsonArray();
        foreach (var c in kat)
        {
            var rows = c!["rows"]!.AsArray().Select(r => r!.AsArray().Select(x => ToF32(x!)).ToArray()).ToArray();
            var q = c["query"]!.AsArray().Select(x => ToF32(x!)).ToArray();
            var (idx, sc) = Query(rows, q, c["top_k"]!.GetValue<int>(), "kat:" + c["name"]!.GetValue<string>());
            var cosBits = new JsonArray();
            foreach (var row in rows) cosBits.Add(Bits((double)cosine.Invoke(null, new object[] { q, row })!));
            katOut.Add(new JsonObject
            {
                ["name"] = c["name"]!.GetValue<string>(),
                ["expect_idx"] = new JsonArray(idx.Select(i => (JsonNode)JsonValue.Create(i)).ToArray()),
                ["expect_score_bits"] = new JsonArray(sc.Select(s => (JsonNode)JsonValue.Create(Bits(s))).ToArray()),
                ["cosine_bits_all_rows"] = cosBits,
            });
        }
        File.WriteAllText(Path.Combine(gold, "from_dotnet", "kat_v1.json"), katOut.ToJsonString(new JsonSerializerOptions { WriteIndented = true }));

        // ---- seeded corpora: the generator first proves itself on the pinned values, then the reference answers
        var synth = JsonNode.Parse(File.ReadAllText(Path.Combine(gold, "synth_v1.json")))!;
        foreach (var (name, bf) in new[] { ("generator", false), ("generator_bf16", true) })
        {
            var g = synth[name]!;
            var got = SynthRows(g["row0"]!.GetValue<long>(), g["n_rows"]!.GetValue<int>(), g["dim"]!.GetValue<int>(), g["seed"]!.GetValue<ulong>(), bf);
            var want = g["f32_hex"]!.AsArray();
            for (int r = 0; r < got.Length; r++)
                for (int col = 0; col < got[r].Length; col++)
                    if (BitConverter.SingleToUInt32Bits(got[r][col]) != want[r]![col]!.GetValue<uint>())
                    { Console.Error.WriteLine($"generator mismatch at {name}[{r}][{col}]: this tool's restatement of rag_synth_rows is wrong"); return 1; }
        }
        var qOut = new JsonArray();
        foreach (var c in synth["queries"]!.AsArray())
        {
            int n = c!["n"]!.GetValue<int>(), d = c["dim"]!.GetValue<int>(), k = c["top_k"]!.GetValue<int>();
            var corpus = SynthRows(0, n, d, c["corpus_seed"]!.GetValue<ulong>());
            var q = SynthRows(0, 1, d, c["query_seed"]!.GetValue<ulong>())[0];
            corpus[7] = (float[])corpus[42].Clone();            // the adversarial inserts of make_golden.py:synth()
            corpus[n - 3] = (float[])corpus[42].Clone();
            corpus[13] = new float[d];
            corpus[21] = (float[])q.Clone();
            corpus[22] = q.Select(x => -x).ToArray();
            var (idx, sc) = Query(corpus, q, k, $"synth:{n}:{d}:{k}");
            qOut.Add(new JsonObject
            {
                ["n"] = n, ["dim"] = d, ["top_k"] = k,
                ["corpus_seed"] = c["corpus_seed"]!.GetValue<ulong>(), ["query_seed"] = c["query_seed"]!.GetValue<ulong>(),
                ["expect_idx"] = new JsonArray(idx.Select(i => (JsonNode)JsonValue.Create(i)).ToArray()),
                ["expect_score_bits"] = new JsonArray(sc.Select(s => (JsonNode)JsonValue.Create(Bits(s))).ToArray()),
            });
        }
        File.WriteAllText(Path.Combine(gold, "from_dotnet", "synth_v1.json"),
                          new JsonObject { ["queries"] = qOut }.ToJsonString(new JsonSerializerOptions { WriteIndented = true }));
        listener.Stop();
        Console.WriteLine($"wrote {gold}/from_dotnet/kat_v1.json and synth_v1.json from {asm.GetName().Name} {asm.GetName().Version} on .NET {Environment.Version}");
        return 0;
    }
}
